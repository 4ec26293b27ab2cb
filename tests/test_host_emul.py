"""CPU: the __host__ __device__ arithmetic of the CUDA env kernel (csrc/env_math.cuh), compiled for the host by a
tests-only shim (tests/host_emul/env_emul.cu), against the reference golden vectors.  This checks the kernel's math
where there is no GPU; the GPU parity tests (-m gpu) check the kernel itself through the C ABI."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest
import torch

from fly_bproject_b200._lib import EnvCfg
from fly_bproject_b200.tables import make_env_cfg
from tests.helpers import ENV_CASES, ROOT, assert_obs_close, load_golden


@pytest.fixture(scope="module")
def emul():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    out_dir = os.path.join(ROOT, "build", "tests")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "libenv_emul.so")
    src = os.path.join(ROOT, "tests", "host_emul", "env_emul.cu")
    deps = [src, os.path.join(ROOT, "fly_bproject_b200", "csrc", "env_math.cuh"),
            os.path.join(ROOT, "include", "flyb200.h")]
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(d) for d in deps):
        subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-std=c++17", "-shared",
                        "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-o", so, src], check=True)
    return C.CDLL(so)


def _p(t):
    return C.c_void_p(t.data_ptr())


@pytest.mark.parametrize("name,asset,n,full", ENV_CASES)
def test_device_math_on_host_matches_golden(emul, name, asset, n, full):
    g = load_golden(name)
    cfg = make_env_cfg(n, asset, full_actions=full)
    a, o = cfg.num_act, 19 + 3 * cfg.num_act
    scaled = torch.zeros(n, a)
    targets = torch.zeros(n * cfg.num_dof)
    emul.emul_env_scale_actions(C.byref(cfg), _p(g["actions_in"]), _p(scaled), _p(targets))
    torch.testing.assert_close(scaled, g["actions_scaled"], rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(targets, g["dof_targets"], rtol=1e-6, atol=1e-6)

    pot, prev = g["potentials_in"].clone(), torch.zeros(n)
    prog, rst = g["progress_in"].clone(), g["reset_in"].clone()
    obs, rew = torch.zeros(n, o), torch.zeros(n)
    up, head = torch.zeros(n, 3), torch.zeros(n, 3)
    emul.emul_env_post_physics(C.byref(cfg), _p(g["root"]), _p(g["dof_states"]), _p(g["force"]),
                               _p(g["actions_scaled"]), _p(pot), _p(prev), _p(prog), _p(rst), _p(obs), _p(rew),
                               _p(up), _p(head))
    assert torch.equal(rst, g["reset"])                     # bit-exact masks
    assert torch.equal(prog, g["progress"])
    assert_obs_close(obs, g["obs"], a)
    torch.testing.assert_close(rew, g["reward"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(pot, g["potentials"], rtol=1e-6, atol=0)
    torch.testing.assert_close(prev, g["prev_potentials"], rtol=0, atol=0)
    torch.testing.assert_close(up, g["up_vec"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(head, g["heading_vec"], rtol=1e-5, atol=1e-6)
