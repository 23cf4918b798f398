/*
 * tests/shim/gst/video/video.h — DECLARATION-ONLY stand-in for the slice of gst-plugins-base's video library that
 * ceracoder_b200/gst/gstcudah264enc.c uses (see tests/shim/gst/gst.h).  TEST INFRASTRUCTURE: no behaviour.
 */
#ifndef SHIM_GST_VIDEO_H
#define SHIM_GST_VIDEO_H
#include <gst/gst.h>
typedef enum { GST_VIDEO_FORMAT_UNKNOWN, GST_VIDEO_FORMAT_ENCODED, GST_VIDEO_FORMAT_I420, GST_VIDEO_FORMAT_YV12, GST_VIDEO_FORMAT_YUY2,
               GST_VIDEO_FORMAT_UYVY, GST_VIDEO_FORMAT_AYUV, GST_VIDEO_FORMAT_RGBx, GST_VIDEO_FORMAT_BGRx, GST_VIDEO_FORMAT_NV12 = 23 } GstVideoFormat;
typedef struct { GstVideoFormat format; } GstVideoFormatInfo;
typedef struct _GstVideoInfo {
    const GstVideoFormatInfo *finfo;
    gint width, height, fps_n, fps_d;
    gsize offset[4];
    gint stride[4];
} GstVideoInfo;
#define GST_VIDEO_INFO_FORMAT(i) ((i)->finfo->format)
#define GST_VIDEO_INFO_WIDTH(i) ((i)->width)
#define GST_VIDEO_INFO_HEIGHT(i) ((i)->height)
#define GST_VIDEO_INFO_FPS_N(i) ((i)->fps_n)
#define GST_VIDEO_INFO_FPS_D(i) ((i)->fps_d)
#define GST_VIDEO_INFO_N_PLANES(i) (3u)
typedef struct _GstVideoFrame {
    GstVideoInfo info;
    GstBuffer *buffer;
    gpointer data[4];
} GstVideoFrame;
gboolean gst_video_frame_map(GstVideoFrame *frame, GstVideoInfo *info, GstBuffer *buffer, GstMapFlags flags);
void gst_video_frame_unmap(GstVideoFrame *frame);
#define GST_VIDEO_FRAME_N_PLANES(f) (GST_VIDEO_INFO_N_PLANES(&(f)->info))
#define GST_VIDEO_FRAME_PLANE_DATA(f, p) ((f)->data[p])
#define GST_VIDEO_FRAME_PLANE_STRIDE(f, p) ((f)->info.stride[p])
#endif
