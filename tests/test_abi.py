"""CPU-only: the C-ABI library loads and exports every symbol include/relayrl_b200.h declares, its
structs have the layout the header states, and without a GPU it fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "relayrl_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rrl_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    from relayrl_b200 import _lib
    L = _lib.lib()
    syms = declared_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(L, s), "library does not export %s" % s
        assert s in _lib.SIGNATURES, "ctypes binding lacks %s" % s
    assert set(_lib.SIGNATURES) == set(syms)


def test_version_status_and_defaults():
    from relayrl_b200 import _lib
    L = _lib.lib()
    assert L.rrl_abi_version() == 1
    assert L.rrl_status_string(0) == b"ok"
    assert b"capacity" in L.rrl_status_string(-3)
    cfg = _lib.Config()
    L.rrl_config_defaults(C.byref(cfg))
    assert (cfg.obs_dim, cfg.act_dim, cfg.hidden, cfg.train_vf_iters, cfg.nranks) == (4, 2, 128, 80, 1)
    # hyper-parameters are f32-rounded doubles (config_loader.rs:148-159)
    assert cfg.gamma == float(np.float32(0.98)) and cfg.vf_lr == float(np.float32(1e-3))
    assert (cfg.adam_beta1, cfg.adam_beta2, cfg.adam_eps) == (0.9, 0.999, 1e-8)
    assert C.sizeof(_lib.Config) == 10 * 4 + 8 + 7 * 8 + 128
    assert C.sizeof(_lib.EpochStats) == 9 * 8 + 2 * 8 + 4 * 4 + 2 * 4


def test_bad_config_rejected_without_touching_the_gpu():
    from relayrl_b200 import _lib
    L = _lib.lib()
    cfg = _lib.Config()
    L.rrl_config_defaults(C.byref(cfg))
    cfg.hidden = 96  # unsupported by the fused kernel
    h = C.c_void_p()
    assert L.rrl_learner_create(C.byref(cfg), C.byref(h)) == -1
    assert b"hidden=96" in L.rrl_last_error(None)
    b = C.c_size_t()
    assert L.rrl_scan_workspace_bytes(1 << 20, C.byref(b)) == 0 and b.value > 0


def test_no_cpu_fallback():
    """Without a CUDA device the product path must raise, not compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from relayrl_b200 import Learner, RelayRLB200Error
    with pytest.raises(RelayRLB200Error) as e:
        Learner(4, 2, 1024)
    assert e.value.status == -2


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "relayrl-prototype_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"import\s+oracle|from\s+oracle|oracle/|liboracle|learner_oracle|scan_ref", txt), \
                    os.path.join(dp, f)
