"""CPU: libflyb200.so builds for sm_100a, loads, and exports every symbol include/flyb200.h declares (no compute)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from tests.helpers import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "flyb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(flyb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound(built_lib):
    from fly_bproject_b200 import _lib
    syms = declared_symbols()
    assert len(syms) >= 14
    handle = C.CDLL(built_lib)
    for s in syms:
        assert hasattr(handle, s), f"{s} declared in include/flyb200.h but not exported"
        assert s in _lib.SIGNATURES, f"{s} has no ctypes signature in fly_bproject_b200/_lib.py"
    assert sorted(_lib.SIGNATURES) == syms
    lib = _lib.lib()
    assert lib.flyb_abi_version() == 7


def test_library_is_sm100a_only(built_lib):
    out = subprocess.run(["cuobjdump", "-lelf", built_lib], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_struct_sizes_match_header(built_lib, tmp_path):
    """ctypes mirrors of the ABI structs must have the C compiler's sizes."""
    from fly_bproject_b200 import _lib
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "flyb200.h"\nint main(){printf("%zu %zu %zu\\n", '
                   'sizeof(FlybEnvCfg), sizeof(FlybNetLayout), sizeof(FlybPpoCfg));printf("%zu %zu\\n", sizeof(FlybStepTail), sizeof(FlybPeerGroup));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    assert sizes == [C.sizeof(_lib.EnvCfg), C.sizeof(_lib.NetLayout), C.sizeof(_lib.PpoCfg), C.sizeof(_lib.StepTail),
                     C.sizeof(_lib.PeerGroup)]


def test_net_layout_and_argument_errors(built_lib):
    from fly_bproject_b200 import _lib
    lib = _lib.lib()
    lay = _lib.NetLayout()
    assert lib.flyb_net_layout(73, 18, C.byref(lay)) == 0
    assert lay.k0_pad == 80 and lay.total % 4 == 0
    assert list(lay.off)[:3] == [0, 256 * 80, 256 * 80 + 256]
    real = 256 * 73 + 256 + 128 * 256 + 128 + 64 * 128 + 64 + 18 * 64 + 18 + 64 * 128 + 64 + 64 + 1
    assert real == 69587                                   # 69 587 parameters at O=73, A=18
    assert lay.total >= real
    assert lib.flyb_net_layout(0, 18, C.byref(lay)) == -1
    assert b"num_obs" in lib.flyb_last_error()
    # compute entry points reject NULL buffers before touching CUDA
    cfg = _lib.EnvCfg()
    assert lib.flyb_env_pre_physics(C.byref(cfg), *([None] * 10)) == -1


def test_tensor_core_shapes(built_lib):
    """flyb_wpack_bytes (host-side, no GPU needed): both compiled shapes of the tcgen05 path announce themselves — the
    18-joint actuation (O = 73, A = 18) and the 42-joint set of BASELINE configs[3] (O = 145, A = 42) — and shapes beyond
    the wide layout report 0 (-> exact-fp32 CUDA-core kernels)."""
    from fly_bproject_b200 import _lib
    lib = _lib.lib()
    sizes = {}
    for o, a in [(73, 18), (145, 42), (100, 27), (37, 6), (163, 48), (145, 49)]:
        lay = _lib.NetLayout()
        assert lib.flyb_net_layout(o, a, C.byref(lay)) == 0
        sizes[(o, a)] = lib.flyb_wpack_bytes(C.byref(lay))
    assert sizes[(73, 18)] > 0 and sizes[(145, 42)] > sizes[(73, 18)]
    assert sizes[(100, 27)] > 0 and sizes[(37, 6)] > 0           # anything up to O = 160 / A = 48 runs the wide forward
    assert sizes[(163, 48)] == 0 and sizes[(145, 49)] == 0


def test_exchange_and_scratch_buffer_sizes(built_lib):
    """Host-side size functions (no GPU): the peer exchange buffer holds the float slots of the split exchange, 256 bytes
    of control and the 64-bit words of the fused one (include/flyb200.h, FlybPeerGroup); the env scratch holds a counter
    pair, one mask per 64 envs (+ 1) and one tagged count per 128-env tile."""
    from fly_bproject_b200 import _lib
    lib = _lib.lib()
    for total, world in [(69587, 2), (69587, 8), (95000, 4), (1, 1)]:
        pad = lib.flyb_peer_total_pad(total)
        assert pad >= total and pad % 64 == 0 and pad - total < 64
        assert lib.flyb_peer_bytes(total, world) == 2 * world * pad * 4 + 256 + 2 * world * pad * 8
    assert lib.flyb_peer_bytes(69587, 9) == 0 and lib.flyb_peer_bytes(0, 2) == 0
    for n in (1, 64, 100, 16384, 18944, 100003):
        masks, tiles = (n + 63) // 64 + 1, (n + 127) // 128 + 1
        assert lib.flyb_env_scratch_bytes(n) == 16 + 8 * masks + 8 * ((tiles + 1) // 2)


def test_product_path_fails_loudly_without_cuda(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    import fly_bproject_b200 as fb
    with pytest.raises(fb.FlybError):
        fb.Fly(fb.make_args(num_envs=8, sim_device="cpu"))
    with pytest.raises(fb.FlybError):
        fb._lib.ptr(torch.zeros(4))


def test_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "fly_bproject_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "/root/reference" not in text, f
