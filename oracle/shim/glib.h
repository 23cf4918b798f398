/* Minimal stand-in for <glib.h>: the reference's balancer core (src/core/*.c) uses GLib only for MIN/MAX
 * (/root/reference/src/core/bitrate_control.c:21, balancer_aimd.c:33).  Test infrastructure only. */
#ifndef CC_SHIM_GLIB_H
#define CC_SHIM_GLIB_H
#include <stddef.h>
#ifndef MIN
#define MIN(a, b) (((a) < (b)) ? (a) : (b))
#endif
#ifndef MAX
#define MAX(a, b) (((a) > (b)) ? (a) : (b))
#endif
#endif
