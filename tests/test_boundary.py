"""The drop-in boundary: every symbol include/ceracuda.h declares is exported by libceracuda.so, the element-shaped
Python mirror keeps the reference's property contract (/root/reference/src/gst/encoder_control.c:23-57), and the
product fails loudly without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess
import pytest
from oracle_lib import ROOT

LIB = os.path.join(ROOT, "ceracoder_b200", "libceracuda.so")


@pytest.fixture(scope="module")
def built():
    subprocess.check_call(["make", "-s", "-C", ROOT, "lib"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return C.CDLL(LIB)


def test_all_declared_symbols_exported(built):
    hdr = open(os.path.join(ROOT, "include", "ceracuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(cc_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 25
    for n in sorted(names):
        assert hasattr(built, n), f"{n} declared in ceracuda.h but not exported"


def test_no_cpu_fallback_without_device(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from ceracoder_b200 import CudaH264Enc, CeraCudaError
    with pytest.raises(CeraCudaError, match="CUDA"):
        CudaH264Enc(64, 48)
    out = (C.c_int32 * 1)()
    a = (C.c_uint8 * 16)()
    assert built.cc_prim_satd4x4(a, a, 1, out) == -2  # CC_ERR_CUDA


def test_product_never_links_the_oracle():
    """The shipped library and package must not link, load, import or name-as-a-path anything under oracle/ or the emulator.
    Comments and docstrings may cite the oracle (file:line references); code may not."""
    out = subprocess.run(["ldd", LIB], capture_output=True, text=True).stdout
    assert "oracle" not in out and "emu" not in out
    needed = subprocess.run(["readelf", "-d", LIB], capture_output=True, text=True).stdout
    assert "ceraoracle" not in needed and "ceracuda_emu" not in needed
    syms = subprocess.run(["nm", "-D", "--undefined-only", LIB], capture_output=True, text=True).stdout
    assert not re.search(r"\bref_[a-z0-9_]+", syms), "libceracuda.so imports oracle symbols (ref_*)"
    import ast
    import io
    import tokenize
    pkg = os.path.join(ROOT, "ceracoder_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            path = os.path.join(dirpath, f)
            if f.endswith(".py"):
                src = open(path).read()
                for node in ast.walk(ast.parse(src)):  # imports
                    mods = [a.name for a in node.names] if isinstance(node, ast.Import) else ([node.module or ""] if isinstance(node, ast.ImportFrom) else [])
                    for mod in mods:
                        assert not re.search(r"oracle|emu_lib|tests", mod), f"{path} imports {mod}"
                code = " ".join(tok.string for tok in tokenize.generate_tokens(io.StringIO(src).readline)
                                if tok.type not in (tokenize.COMMENT, tokenize.STRING))  # identifiers outside comments / docstrings
                assert not re.search(r"oracle|libceraoracle|emu_lib", code, re.I), f"{path} names the oracle in code"
                strings = [tok.string for tok in tokenize.generate_tokens(io.StringIO(src).readline) if tok.type == tokenize.STRING]
                assert not any(re.search(r"libceraoracle|oracle/|oracle_lib|_emu\.so", s) and not s.lstrip("rRbBuU").startswith(('\"\"\"', "\'\'\'")) for s in strings), \
                    f"{path} carries an oracle / emulator path in a string literal"
            elif f.endswith((".c", ".cu", ".cuh", ".h")):
                src = re.sub(r"/\*.*?\*/|//[^\n]*", "", open(path).read(), flags=re.S)  # code without comments
                assert not re.search(r"#include\s*[<\"][^>\"]*(oracle|ref_h264|ref_int|ref_tables)", src), f"{path} includes an oracle header"
                assert not re.search(r"\bref_[a-z0-9_]+\s*\(", src), f"{path} calls an oracle function"
                assert not re.search(r"dlopen|libceraoracle", src), f"{path} loads a library at run time"


def test_the_oracle_check_catches_a_regression(tmp_path):
    """the scan above is not vacuous: a file that imports the oracle binding, or calls ref_*, fails the same patterns"""
    import ast
    bad_py = "from oracle_lib import OracleEncoder\n"
    assert any(re.search(r"oracle|emu_lib|tests", n.module or "") for n in ast.walk(ast.parse(bad_py)) if isinstance(n, ast.ImportFrom))
    bad_c = "int f(void) { return ref_satd_4x4(a, 4, b, 4); } /* comment: ref_ok( */"
    assert re.search(r"\bref_[a-z0-9_]+\s*\(", re.sub(r"/\*.*?\*/|//[^\n]*", "", bad_c, flags=re.S))
    assert not re.search(r"\bref_[a-z0-9_]+\s*\(", re.sub(r"/\*.*?\*/|//[^\n]*", "", "/* ref_ok( */ int g;", flags=re.S))


def test_property_contract_on_emulated_element():
    from emu_lib import emu_library
    from ceracoder_b200 import CudaH264Enc
    L = emu_library()
    e = CudaH264Enc(64, 48, bitrate=6000, speed_preset="veryfast", key_int_max=60, name="venc_kbps", lib=L)
    assert e.get_property("bitrate") == 6000
    e.set_property("bps", 2500)           # what encoder_control.c:53 writes for a venc_kbps element: bps/1000
    assert e.get_property("bitrate") == 2500
    e.set_property("bitrate", 12000)      # what the docs promise (README.md:256)
    assert e.get_property("bps") == 12000
    e.set_property("bitrate", 10)         # clamped to the reference's legal range (bitrate_control.h:30-31)
    assert e.get_property("bitrate") == 300
    e.set_property("bitrate", 10 ** 6)
    assert e.get_property("bitrate") == 30000
    with pytest.raises(AttributeError):
        e.set_property("qp-min", 3)
    e.close()
    e2 = CudaH264Enc(64, 48, name="venc_bps", lib=L)
    e2.set_property("bps", 4200000)       # venc_bps elements get bit/s (encoder_control.c:29-30)
    assert e2.get_property("bitrate") == 4200
    e2.close()
    with pytest.raises(ValueError):
        CudaH264Enc(64, 48, name="x264enc0", lib=L)
    with pytest.raises(ValueError):
        CudaH264Enc(64, 48, speed_preset="warp9", lib=L)


def test_parameter_errors_on_emulated_library():
    from emu_lib import emu_library
    from ceracoder_b200 import CudaH264Enc, CeraCudaError
    L = emu_library()
    for w, h in ((8, 48), (64, 47), (65, 48), (20000, 48)):
        with pytest.raises(CeraCudaError, match="invalid parameter"):
            CudaH264Enc(w, h, lib=L)
    assert L.cc_strerror(-5).decode().startswith("output buffer")
    assert b"ceracuda" in L.cc_version()


def test_stage_names_of_the_mirror_match_the_header():
    """CudaH264Enc.STAGES names the CC_STAGE_* enum of include/ceracuda.h in order (bench.py's stage_ms_per_frame keys)"""
    from ceracoder_b200.encoder import CudaH264Enc
    hdr = open(os.path.join(ROOT, "include", "ceracuda.h")).read()
    m = re.search(r"enum \{ (CC_STAGE_UPLOAD.*?CC_STAGE_COUNT) \}", hdr, re.S)
    names = [n.strip()[len("CC_STAGE_"):].lower() for n in m.group(1).replace("= 0", "").split(",")]
    assert names[-1] == "count" and tuple(names[:-1]) == CudaH264Enc.STAGES
