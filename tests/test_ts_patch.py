"""integration/ts-cuda-encoder.patch adds `encoder: "x264" | "cuda"` to the reference's TypeScript pipeline builder
(/root/reference/bindings/typescript/src/pipeline/generic-builder.ts:43-55).  No node / bun on this image, so the patch is applied
to a scratch copy of the reference files and the generated encoder line is evaluated from the patched template text; the string
checks mirror /root/reference/bindings/typescript/src/pipeline/index.test.ts:144-163."""
import os
import re
import shutil
import subprocess
import pytest
from oracle_lib import ROOT

REF = "/root/reference/bindings/typescript/src/pipeline"
PATCH = os.path.join(ROOT, "integration", "ts-cuda-encoder.patch")


def test_patch_is_committed_and_touches_only_the_pipeline_builder():
    txt = open(PATCH).read()
    files = re.findall(r"^\+\+\+ (\S+)", txt, re.M)
    assert files == ["bindings/typescript/src/pipeline/types.ts", "bindings/typescript/src/pipeline/generic-builder.ts",
                     "bindings/typescript/src/pipeline/index.test.ts"]
    assert "name=venc_kbps" in txt and "cudah264enc" in txt


@pytest.mark.skipif(not os.path.isdir(REF), reason="the reference tree only exists in the build container")
def test_patch_applies_and_generates_the_cuda_encoder_line(tmp_path):
    dst = tmp_path / "bindings" / "typescript" / "src" / "pipeline"
    dst.mkdir(parents=True)
    for f in ("types.ts", "generic-builder.ts", "index.test.ts", "index.ts", "common.ts"):
        shutil.copy(os.path.join(REF, f), dst / f)
    r = subprocess.run(["patch", "-p0", "--no-backup-if-mismatch", "-i", PATCH], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    gb = (dst / "generic-builder.ts").read_text()
    factory = dict(re.findall(r"^\t(\w+): \"(\w+)\",$", re.search(r"ENCODER_FACTORY[^{]*\{(.*?)\};", gb, re.S).group(1), re.M))
    presets = {k: int(v) for k, v in re.findall(r"^\t(\w+): (\d+),$", re.search(r"X264_PRESET_MAP[^{]*\{(.*?)\};", gb, re.S).group(1), re.M)}
    tmpl = re.search(r"return `(\$\{factory\}[^`]*)`;", gb).group(1)

    def line(encoder=None, preset="superfast", framerate=30):
        gop = round(framerate * 2)  # calculateGop, common.ts:104-107
        return tmpl.replace("${factory}", factory[encoder or "x264"]).replace("${presetValue}", str(presets[preset])).replace("${gop}", str(gop))

    assert factory == {"x264": "x264enc", "cuda": "cudah264enc"}
    assert line().startswith("x264enc speed-preset=2 key-int-max=60 name=venc_kbps ! h264parse config-interval=-1")  # default unchanged
    cuda = line("cuda", "veryfast", 60)
    assert cuda.startswith("cudah264enc speed-preset=3 key-int-max=120 name=venc_kbps ! h264parse config-interval=-1")
    # the builder's validator (index.ts:139-149) keeps passing: it only looks for the element names
    assert "name=venc_kbps" in cuda and "mux" in cuda
    types = (dst / "types.ts").read_text()
    assert 'softwareEncoderSchema = z.enum(["x264", "cuda"])' in types and "encoder: softwareEncoderSchema.optional()" in types
    tests = (dst / "index.test.ts").read_text()
    assert tests.count('overrides: { encoder: "cuda"') == 2
    # and the generated line is one the element accepts (same check as for the committed pipeline files)
    from test_element_source import _pspecs, _encoder_props, _accepts, SRC
    specs = _pspecs(SRC)
    props = _encoder_props(cuda)
    props.pop("name")
    assert all(_accepts(specs[k], v) for k, v in props.items())
