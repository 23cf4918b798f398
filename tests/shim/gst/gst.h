/*
 * tests/shim/gst/gst.h — DECLARATION-ONLY stand-in for the slice of GLib / GObject / GStreamer core that
 * ceracoder_b200/gst/gstcudah264enc.c uses.  TEST INFRASTRUCTURE: types, macros and prototypes, no behaviour, nothing links
 * against it.  Neither this image nor the GPU box has GLib or GStreamer headers; `gcc -fsyntax-only -Wall -Werror -Itests/shim`
 * over the element source (tests/test_element_source.py) is what keeps that file honest about argument counts, types and
 * missing declarations until a box with the real headers builds it (`make gst`).  Signatures follow the GLib 2.56 / GStreamer
 * 1.14 API documentation (the versions the reference asks for, /root/reference/docs/dependencies.md:12).
 */
#ifndef SHIM_GST_H
#define SHIM_GST_H
#include <stddef.h>
#include <stdint.h>
#include <limits.h>

/* ---- GLib ------------------------------------------------------------------------------------------- */
typedef int gint;
typedef unsigned int guint;
typedef int gboolean;
typedef char gchar;
typedef unsigned char guint8;
typedef size_t gsize;
typedef void *gpointer;
typedef const void *gconstpointer;
typedef uint64_t guint64;
typedef int64_t gint64;
typedef size_t GType;
#ifndef TRUE
#define TRUE 1
#define FALSE 0
#endif
#define G_MAXINT INT_MAX
#define G_GNUC_UNUSED __attribute__((unused))
void g_free(gpointer mem);
gboolean g_str_has_prefix(const gchar *str, const gchar *prefix);
void g_atomic_int_set(volatile gint *atomic, gint newval);
gint g_atomic_int_get(const volatile gint *atomic);
gboolean g_once_init_enter(volatile void *location);
void g_once_init_leave(volatile void *location, gsize result);

/* ---- GObject ------------------------------------------------------------------------------------------ */
typedef struct _GObject GObject;
typedef struct _GObjectClass GObjectClass;
typedef struct _GValue GValue;
typedef struct _GParamSpec GParamSpec;
typedef struct _GTypeInstance { gpointer g_class; } GTypeInstance;
struct _GObject { GTypeInstance g_type_instance; guint ref_count; gpointer qdata; };
struct _GObjectClass {
    GType g_type;
    void (*set_property)(GObject *object, guint property_id, const GValue *value, GParamSpec *pspec);
    void (*get_property)(GObject *object, guint property_id, GValue *value, GParamSpec *pspec);
    void (*dispose)(GObject *object);
    void (*finalize)(GObject *object);
};
typedef enum { G_PARAM_READABLE = 1 << 0, G_PARAM_WRITABLE = 1 << 1, G_PARAM_READWRITE = 3, G_PARAM_STATIC_STRINGS = 0xe0 } GParamFlags;
typedef struct { gint value; const gchar *value_name; const gchar *value_nick; } GEnumValue;
typedef struct { guint value; const gchar *value_name; const gchar *value_nick; } GFlagsValue;
#define G_TYPE_STRING ((GType)(16 << 2))
GType g_enum_register_static(const gchar *name, const GEnumValue *const_static_values);
GType g_flags_register_static(const gchar *name, const GFlagsValue *const_static_values);
guint g_value_get_uint(const GValue *value);
gint g_value_get_int(const GValue *value);
gboolean g_value_get_boolean(const GValue *value);
gint g_value_get_enum(const GValue *value);
guint g_value_get_flags(const GValue *value);
void g_value_set_uint(GValue *value, guint v);
void g_value_set_int(GValue *value, gint v);
void g_value_set_boolean(GValue *value, gboolean v);
void g_value_set_enum(GValue *value, gint v);
void g_value_set_flags(GValue *value, guint v);
GParamSpec *g_param_spec_uint(const gchar *name, const gchar *nick, const gchar *blurb, guint minimum, guint maximum, guint default_value, GParamFlags flags);
GParamSpec *g_param_spec_int(const gchar *name, const gchar *nick, const gchar *blurb, gint minimum, gint maximum, gint default_value, GParamFlags flags);
GParamSpec *g_param_spec_boolean(const gchar *name, const gchar *nick, const gchar *blurb, gboolean default_value, GParamFlags flags);
GParamSpec *g_param_spec_enum(const gchar *name, const gchar *nick, const gchar *blurb, GType enum_type, gint default_value, GParamFlags flags);
GParamSpec *g_param_spec_flags(const gchar *name, const gchar *nick, const gchar *blurb, GType flags_type, guint default_value, GParamFlags flags);
void g_object_class_install_property(GObjectClass *oclass, guint property_id, GParamSpec *pspec);
void shim_warn_invalid_property_id(GObject *object, guint property_id, GParamSpec *pspec);
#define G_OBJECT_WARN_INVALID_PROPERTY_ID(object, property_id, pspec) shim_warn_invalid_property_id((GObject *)(object), (property_id), (pspec))
#define G_OBJECT_CLASS(klass) ((GObjectClass *)(klass))

/* G_DECLARE_FINAL_TYPE / G_DEFINE_TYPE: the declarations the real macros expand to that the element relies on */
#define G_DECLARE_FINAL_TYPE(ModuleObjName, module_obj_name, MODULE, OBJ_NAME, ParentName) \
    GType module_obj_name##_get_type(void);                                                \
    typedef struct _##ModuleObjName ModuleObjName;                                         \
    typedef struct { ParentName##Class parent_class; } ModuleObjName##Class;               \
    static inline ModuleObjName *MODULE##_##OBJ_NAME(gpointer ptr) { return (ModuleObjName *)ptr; }
#define G_DEFINE_TYPE(TypeName, type_name, TYPE_PARENT)                          \
    static void type_name##_init(TypeName *self);                                \
    static void type_name##_class_init(TypeName##Class *klass);                  \
    static gpointer type_name##_parent_class = NULL;                             \
    GType type_name##_get_type(void) {                                           \
        (void)type_name##_init; (void)type_name##_class_init; (void)TYPE_PARENT; \
        return (GType)0;                                                         \
    }

/* ---- GStreamer core ------------------------------------------------------------------------------------ */
typedef struct _GstObject { GObject object; gpointer lock; gchar *name; } GstObject;
typedef struct _GstObjectClass { GObjectClass parent_class; } GstObjectClass;
typedef struct _GstElement { GstObject object; } GstElement;
typedef struct _GstElementClass { GstObjectClass parent_class; } GstElementClass;
typedef struct _GstPlugin GstPlugin;
typedef struct _GstCaps GstCaps;
typedef struct _GstBuffer GstBuffer;
typedef struct _GstAllocator GstAllocator;
typedef struct _GstAllocationParams GstAllocationParams;
typedef struct _GstDebugCategory GstDebugCategory;
typedef guint64 GstClockTime;
typedef enum { GST_FLOW_OK = 0, GST_FLOW_NOT_NEGOTIATED = -4, GST_FLOW_ERROR = -5 } GstFlowReturn;
typedef enum { GST_PAD_UNKNOWN, GST_PAD_SRC, GST_PAD_SINK } GstPadDirection;
typedef enum { GST_PAD_ALWAYS, GST_PAD_SOMETIMES, GST_PAD_REQUEST } GstPadPresence;
typedef enum { GST_MAP_READ = 1, GST_MAP_WRITE = 2 } GstMapFlags;
typedef enum { GST_RANK_NONE = 0, GST_RANK_MARGINAL = 64, GST_RANK_SECONDARY = 128, GST_RANK_PRIMARY = 256 } GstRank;
typedef struct { GstCaps *caps; const char *string; } GstStaticCaps;
typedef struct { const gchar *name_template; GstPadDirection direction; GstPadPresence presence; GstStaticCaps static_caps; } GstStaticPadTemplate;
#define GST_STATIC_CAPS(string) { NULL, string }
#define GST_STATIC_PAD_TEMPLATE(padname, dir, pres, caps) { padname, dir, pres, caps }
#define GST_PARAM_MUTABLE_PLAYING (1 << 12)
#define GST_OBJECT(obj) ((GstObject *)(obj))
#define GST_ELEMENT_CLASS(klass) ((GstElementClass *)(klass))
void shim_object_lock(GstObject *obj);
void shim_object_unlock(GstObject *obj);
#define GST_OBJECT_LOCK(obj) shim_object_lock(GST_OBJECT(obj))
#define GST_OBJECT_UNLOCK(obj) shim_object_unlock(GST_OBJECT(obj))
gchar *gst_object_get_name(GstObject *object);
GstCaps *gst_caps_new_simple(const char *media_type, const char *fieldname, ...);
GstBuffer *gst_buffer_new_allocate(GstAllocator *allocator, gsize size, GstAllocationParams *params);
gsize gst_buffer_fill(GstBuffer *buffer, gsize offset, gconstpointer src, gsize size);
void gst_element_class_set_static_metadata(GstElementClass *klass, const gchar *longname, const gchar *classification, const gchar *description, const gchar *author);
void gst_element_class_add_static_pad_template(GstElementClass *klass, GstStaticPadTemplate *static_templ);
gboolean gst_element_register(GstPlugin *plugin, const gchar *name, guint rank, GType type);
#define GST_DEBUG_CATEGORY_STATIC(cat) static GstDebugCategory *cat G_GNUC_UNUSED = NULL
#define GST_DEBUG_CATEGORY_INIT(cat, name, color, description) do { (void)(name); (void)(color); (void)(description); (cat) = NULL; } while (0)
/* GST_ELEMENT_ERROR(el, domain, code, (text...), (debug...)): both message arguments are parenthesised printf lists or (NULL) */
void shim_element_error(gpointer element, const char *domain, const char *code, gchar *text, gchar *debug);
gchar *shim_error_printf(const gchar *format, ...) __attribute__((format(printf, 1, 2)));
#define SHIM_ERR_ARGS_NULL NULL
#define GST_ELEMENT_ERROR(el, domain, code, text, debug) shim_element_error((el), #domain, #code, shim_error_printf text, shim_error_printf debug)
#define GST_VERSION_MAJOR 1
#define GST_VERSION_MINOR 14
#define GST_PLUGIN_DEFINE(major, minor, name, description, init, version, license, package, origin) \
    gboolean (*const shim_plugin_init_##name)(GstPlugin *) G_GNUC_UNUSED = init;                      \
    const char *const shim_plugin_desc_##name[] G_GNUC_UNUSED = { description, version, license, package, origin, PACKAGE };
#endif
