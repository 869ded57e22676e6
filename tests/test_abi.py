"""CPU: the C-ABI library loads, exports every symbol include/radiomedium_b200.h declares, its struct
layouts match the ctypes mirror, and -- there being no GPU here -- refuses to create a handle
instead of falling back to a CPU path."""
import ctypes as C
import os
import re
import subprocess
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "radiomedium_b200.h")


@pytest.fixture(scope="module")
def rsb():
    import __graft_entry__ as g
    g.build()
    import radio_sim_b200
    return radio_sim_b200


def declared_symbols():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"\b(rmb_[a-z0-9_]+)\s*\(", text)))


def test_header_and_ctypes_symbol_lists_agree(rsb):
    assert declared_symbols() == sorted(rsb.SYMBOLS)


def test_library_exports_every_declared_symbol(rsb):
    lib = rsb.load()
    out = subprocess.run(["nm", "-D", "--defined-only", rsb.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (rmb_\w+)", out))
    for s in declared_symbols():
        assert s in exported and hasattr(lib, s), s
    assert lib.rmb_version() == 2


def test_struct_layouts_match_a_c_compiler(rsb, tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "radiomedium_b200.h"\nint main(void){printf("%zu %zu %zu %zu\\n",'
                   'sizeof(rmb_config),sizeof(rmb_tx_record),sizeof(rmb_result),sizeof(rmb_timing));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    sizes = [int(v) for v in subprocess.check_output([str(exe)], text=True).split()]
    from radio_sim_b200 import _abi as A
    assert sizes == [C.sizeof(A.Config), C.sizeof(A.TxRecord), C.sizeof(A.Result), C.sizeof(A.Timing)]


def test_config_defaults_follow_the_reference(rsb):
    from radio_sim_b200 import _abi as A
    cfg = A.Config()
    rsb.load().rmb_config_default(C.byref(cfg), A.MODEL_UDGM)
    assert cfg.struct_size == C.sizeof(A.Config)
    assert (cfg.success_ratio_tx, cfg.success_ratio_rx, cfg.transmission_range) == (1.0, 1.0, 50.0)   # UDGMRadioMedium.java:17-23
    assert cfg.constant_range == 100.0 and cfg.noise_floor_dbm == -100.0


def test_no_gpu_means_error_not_fallback(rsb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(rsb.MediumError) as e:
        rsb.Medium("sinr")
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "radio-sim_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".java", ".cpp")):
                text = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle_py" not in text and "liboracle" not in text and "medium_oracle" not in text, f
