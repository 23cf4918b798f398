"""The outer boundary as source: ceracoder_b200/gst/gstcudah264enc.c is held to the GStreamer API by compiling it against the
declaration-only headers under tests/shim (this image has no GLib / GStreamer), and its property table is held to every
pipeline file that names it — gst_parse_launch (/root/reference/src/io/pipeline_loader.c:57-72) fails on a property the
element does not have or a value its pspec does not accept."""
import os
import re
import subprocess
import pytest
from oracle_lib import ROOT

SRC = os.path.join(ROOT, "ceracoder_b200", "gst", "gstcudah264enc.c")
GCC = ["gcc", "-fsyntax-only", "-std=gnu11", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "tests", "shim"), "-I" + os.path.join(ROOT, "include")]


def test_element_compiles_against_the_declaration_shim():
    r = subprocess.run(GCC + [SRC], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


@pytest.mark.parametrize("snippet", [
    "GstFlowReturn f(GstVideoEncoder *e) { return gst_video_encoder_finish_frame(e); }",                 # argument count
    "void f(GstVideoCodecFrame *fr) { fr->output = 0; }",                                                # unknown member
    "void f(GObjectClass *oc) { g_object_class_install_property(oc, 1, g_param_spec_enum(\"a\", \"b\", \"c\", 3, 0)); }",  # missing type argument
    "void f(GstVideoEncoder *e) { gst_video_encoder_drop_it(e); }",                                      # undeclared function
])
def test_shim_rejects_api_misuse(snippet, tmp_path):
    """negative control: the shim is strict enough to fail on the classes of mistake an uncompiled element can carry"""
    p = tmp_path / "bad.c"
    p.write_text('#include <gst/video/gstvideoencoder.h>\n' + snippet + "\n")
    assert subprocess.run(GCC + [str(p)], capture_output=True).returncode != 0


def _pspecs(src):
    """{property: (kind, min, max | nick table)} from the g_param_spec_* calls of class_init"""
    txt = open(src).read()
    tables = {}
    for m in re.finditer(r"static const G(Enum|Flags)Value values\[\] = \{(.*?)\};\s*if \(g_once_init_enter.*?g_(?:enum|flags)_register_static\(\"(\w+)\"", txt, re.S):
        tables[m.group(3)] = {nick: int(val, 0) for val, nick in re.findall(r"\{(0x[0-9a-f]+|\d+), \"[^\"]*\", \"([^\"]+)\"\}", m.group(2))}
    gtype = dict(re.findall(r"#define (GST_TYPE_CUDA_H264_ENC_\w+) \((\w+)_get_type\(\)\)", txt))
    regname = {fn: name for fn, name in re.findall(r"static GType (\w+)_get_type\(void\).*?register_static\(\"(\w+)\"", txt, re.S)}
    out = {}
    for kind, name, rest in re.findall(r"g_param_spec_(uint|int|boolean|enum|flags)\(\"([a-z-]+)\", \"[^\"]*\", \"[^\"]*\", ([^;]*?), rwp?\)\);", txt):
        args = [a.strip() for a in rest.split(",")]
        if kind in ("uint", "int"):
            out[name] = (kind, int(args[0], 0), (2 ** 31 - 1) if args[1] == "G_MAXINT" else int(args[1], 0))
        elif kind == "boolean":
            out[name] = (kind, None, None)
        else:
            out[name] = (kind, tables[regname[gtype[args[0]]]], int(args[1], 0))
    return out


def _accepts(spec, value):
    kind = spec[0]
    if kind in ("uint", "int"):
        return re.fullmatch(r"-?\d+", value) is not None and spec[1] <= int(value) <= spec[2]
    if kind == "boolean":
        return value.lower() in ("true", "false", "0", "1", "yes", "no")
    nicks = spec[1]
    if kind == "enum":   # gst_value_deserialize: by value or by nick
        return value in nicks or (value.isdigit() and int(value) in nicks.values())
    return all(v in nicks or (v.isdigit() and int(v) <= sum(nicks.values())) for v in value.split("+"))  # flags: nick+nick or a number


def _encoder_props(line):
    m = re.search(r"cudah264enc((?:\s+[a-z-]+=\S+)*)", line)
    assert m, line
    return dict(kv.split("=", 1) for kv in m.group(1).split())


def test_every_pipeline_line_parses_against_the_property_table():
    specs = _pspecs(SRC)
    assert {"bitrate", "bps", "speed-preset", "key-int-max", "tune", "byte-stream", "threads"} <= set(specs)
    d = os.path.join(ROOT, "pipeline", "generic")
    files = sorted(f for f in os.listdir(d) if f.startswith("cuda_h264_"))
    assert len(files) >= 10
    for f in files:
        props = _encoder_props(open(os.path.join(d, f)).read())
        assert props.pop("name") in ("venc_kbps", "venc_bps")  # /root/reference/src/gst/encoder_control.c:29-32
        for k, v in props.items():
            assert k in specs, f"{f}: cudah264enc has no property {k!r}"
            assert _accepts(specs[k], v), f"{f}: {k}={v} is outside the pspec {specs[k]}"


def test_baseline_spelling_and_runtime_writes_parse():
    """BASELINE.json spells the x264enc line with nicks (`speed-preset=veryfast tune=zerolatency`); the reference's files use the
    number; ceracoder writes `bps` as a C int (/root/reference/src/gst/encoder_control.c:53) in kbit/s or bit/s"""
    specs = _pspecs(SRC)
    props = _encoder_props("cudah264enc speed-preset=veryfast tune=zerolatency bitrate=6000 key-int-max=60 byte-stream=true threads=4 name=venc_kbps")
    props.pop("name")
    for k, v in props.items():
        assert _accepts(specs[k], v), (k, v)
    assert specs["speed-preset"][1]["veryfast"] == 3 and specs["speed-preset"][1]["superfast"] == 2   # x264enc's numbering
    assert specs["tune"][1]["zerolatency"] == 0x4
    assert not _accepts(specs["speed-preset"], "warp9") and not _accepts(specs["tune"], "psnr")
    for bps in (300, 30000, 300000, 30000000):  # legal range in both units (/root/reference/src/core/bitrate_control.h:30-31)
        assert _accepts(specs["bps"], str(bps))
    assert _accepts(specs["bitrate"], "300") and _accepts(specs["bitrate"], "30000") and not _accepts(specs["bitrate"], "30001")


def test_advertised_profile_is_the_profile_of_the_sps():
    txt = open(SRC).read()
    profiles = set(re.findall(r"profile=\(string\)([a-z-]+)", txt)) | set(re.findall(r"\"profile\", G_TYPE_STRING, \"([a-z-]+)\"", txt))
    bs = open(os.path.join(ROOT, "ceracoder_b200", "csrc", "cc_bitstream.c")).read()
    idc = int(re.search(r"cc_bw_put\(&b, 8, (\d+)\);\s*/\* profile_idc", bs).group(1))
    assert profiles == {{66: "constrained-baseline", 77: "main", 100: "high"}[idc]}


def test_encoder_handle_is_only_touched_under_the_object_lock():
    """set_property runs on the main-loop thread while set_format swaps the handle on the streaming thread"""
    txt = open(SRC).read()
    for fn in ("set_bitrate", "close_codec", "gst_cuda_h264_enc_set_format"):
        body = re.search(r"\b" + fn + r"\([^)]*\) \{(.*?)\n\}", txt, re.S).group(1)
        depth = 0
        for tok in re.finditer(r"GST_OBJECT_LOCK|GST_OBJECT_UNLOCK|self->enc\b", body):
            if tok.group() == "GST_OBJECT_LOCK":
                depth += 1
            elif tok.group() == "GST_OBJECT_UNLOCK":
                depth -= 1
            else:
                assert depth == 1, f"{fn}: self->enc used outside GST_OBJECT_LOCK"
        assert depth == 0
    hf = re.search(r"gst_cuda_h264_enc_handle_frame\([^)]*\) \{(.*?)\n\}", txt, re.S).group(1)
    assert "gst_video_codec_frame_unref" not in hf  # every exit goes through finish_frame
    assert hf.count("gst_video_encoder_finish_frame") == 4
