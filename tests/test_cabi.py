"""CPU: the C-ABI library loads, exports what include/cardioflow.h declares, and the host mirror of the
reference's module API validates arguments / keeps the state-dict contract.  No compute calls here."""
import ctypes

import numpy as np
import pytest
import torch

import cardioflow_b200 as cf
from cardioflow_b200 import _cabi


def test_library_loads_and_exports_every_declared_symbol():
    lib = _cabi.lib()
    declared = _cabi.declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/cardioflow.h but not exported"
        assert name in _cabi._SIGNATURES, f"{name} has no ctypes signature"
    assert lib.cf_version() >= 100


def test_pyramid_geometry():
    lib = _cabi.lib()
    B, H, W, L = 2, 24, 24, 4
    sizes = [24, 12, 6, 3]
    off = 0
    for lvl, s in enumerate(sizes):
        assert lib.cf_corr_pyramid_level_offset(B, H, W, lvl) == off
        off += -(-(B * H * W * s * s * 4) // 256) * 256
        h, w = ctypes.c_int(), ctypes.c_int()
        assert lib.cf_pyramid_level_dims(H, W, lvl, ctypes.byref(h), ctypes.byref(w)) == 0
        assert (h.value, w.value) == (s, s)
    assert lib.cf_corr_pyramid_bytes(B, H, W, L) == off


def test_argument_errors_are_reported_without_a_gpu():
    lib = _cabi.lib()
    assert lib.cf_corr_lookup(None, None, 1, 16, 16, 4, 4, None, None) == _cabi.CF_ERR_BAD_ARG
    assert b"null" in lib.cf_last_error()
    assert lib.cf_corr_lookup_taps(ctypes.c_void_p(256), 1, 8, 8, 4, 4, ctypes.c_void_p(256), ctypes.c_void_p(256), None) \
        == _cabi.CF_ERR_UNSUPPORTED                      # 8x8 too small for 4 levels (raft.py:368-383)
    assert lib.cf_corr_lookup_taps(ctypes.c_void_p(256), 1, 16, 16, 4, 5, ctypes.c_void_p(256), ctypes.c_void_p(256), None) \
        == _cabi.CF_ERR_UNSUPPORTED                      # radius > CF_MAX_RADIUS


def test_no_cpu_fallback():
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        cf.CorrBlock(torch.randn(1, 8, 16, 16), torch.randn(1, 8, 16, 16))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        cf.SpatialTransformer((8, 8))(torch.zeros(1, 2, 8, 8), torch.zeros(1, 1, 8, 8))
    with pytest.raises(RuntimeError):
        cf.coords_grid(1, 4, 4, device="cpu")


def test_corrblock_validation_matches_torchvision_errors():
    with pytest.raises(ValueError, match="same shape"):
        cf.CorrBlock(torch.randn(1, 8, 16, 16), torch.randn(1, 8, 16, 24))
    with pytest.raises(ValueError, match="too small"):
        cf.CorrBlock(torch.randn(1, 8, 8, 8), torch.randn(1, 8, 8, 8))


def test_spatial_transformer_module_contract(golden_warp):
    st = cf.SpatialTransformer((6, 5))
    sd = st.state_dict()
    assert list(sd.keys()) == ["grid"] and sd["grid"].dtype == torch.float32
    assert np.array_equal(sd["grid"].numpy(), golden_warp["state.grid_6x5"])
    # strict load of a state dict produced by the REFERENCE module (frozen in the fixture)
    st.load_state_dict({"grid": torch.from_numpy(golden_warp["state.grid_6x5"])}, strict=True)
    import inspect
    assert list(inspect.signature(st.forward).parameters) == ["flow", "original", "mode"]
    with pytest.raises(ValueError):
        cf.SpatialTransformer((4, 4, 4))
    v = cf.VecInt((8, 8), 3)
    assert list(v.state_dict().keys()) == ["transformer.grid"] and v.scale == 0.125


def test_shim_import_hook_against_a_regular_package_tree(tmp_path):
    """The zero-change drop-in: with shim/ on PYTHONPATH (sitecustomize installs the import hook) the reference's OWN import
    lines resolve to cardioflow_b200 although nnunet / nnunet.lib / nnunet.network_architecture are regular packages with
    __init__.py files, and BasicUpdateBlock still comes from the tree's own raft_initial.py (SegFlowGaussian.py:35,42)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    tree = tmp_path / "tree"
    for pkg in ("nnunet", "nnunet/lib", "nnunet/network_architecture"):
        (tree / pkg).mkdir(parents=True)
        (tree / pkg / "__init__.py").write_text("")
    (tree / "nnunet/lib/raft_initial.py").write_text(
        "class BasicUpdateBlock: origin = 'tree'\nclass CorrBlock: origin = 'tree'\ndef coords_grid(*a, **k): return 'tree'\n")
    (tree / "nnunet/network_architecture/integration.py").write_text(
        "class SpatialTransformer: origin = 'tree'\nclass VecInt: origin = 'tree'\nEXTRA = 41\n")
    prog = (
        "from nnunet.lib.raft_initial import CorrBlock, BasicUpdateBlock, coords_grid\n"
        "from nnunet.network_architecture.integration import SpatialTransformer, VecInt, SpatialTransformerContour\n"
        "import nnunet.network_architecture.integration as integ, nnunet.lib.raft_initial as ri\n"
        "import cardioflow_b200 as cf\n"
        "assert CorrBlock is cf.CorrBlock and coords_grid is cf.coords_grid, CorrBlock\n"
        "assert SpatialTransformer is cf.SpatialTransformer and VecInt is cf.VecInt\n"
        "assert SpatialTransformerContour is cf.SpatialTransformerContour\n"
        "assert BasicUpdateBlock.origin == 'tree' and integ.EXTRA == 41\n"
        "assert ri.__name__ == 'nnunet.lib.raft_initial' and 'shim' in ri.__file__\n"
        "try:\n    ri.NoSuchThing\nexcept AttributeError: pass\nelse: raise SystemExit('missing names must raise')\n"
        "print('hook ok')\n")
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([os.path.join(root, "shim"), str(tree)]))
    r = subprocess.run([sys.executable, "-c", prog], capture_output=True, text=True, env=env, cwd=str(tmp_path))
    assert r.returncode == 0 and "hook ok" in r.stdout, r.stdout + r.stderr
    # and a tree WITHOUT raft_initial.py (the upstream state): CorrBlock resolves, BasicUpdateBlock raises a clear ImportError
    os.remove(tree / "nnunet/lib/raft_initial.py")
    prog2 = ("from nnunet.lib.raft_initial import CorrBlock, coords_grid\n"
             "try:\n    from nnunet.lib.raft_initial import BasicUpdateBlock\n"
             "except ImportError as e: print('clear error:', e)\nelse: raise SystemExit('expected ImportError')\n")
    r = subprocess.run([sys.executable, "-c", prog2], capture_output=True, text=True, env=env, cwd=str(tmp_path))
    assert r.returncode == 0 and "clear error" in r.stdout, r.stdout + r.stderr


def test_header_is_plain_c_and_a_c_host_links(tmp_path):
    """include/cardioflow.h must be usable from a C host (the drop-in boundary has no C++ or torch types): compile a C
    program with gcc -std=c99 -pedantic against the header, link libcardioflow.so, and make calls that need no GPU."""
    import os
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib_dir = os.path.dirname(_cabi.LIB_PATH)
    _cabi.lib()                                     # fails loudly if the library has not been built
    src = tmp_path / "host.c"
    src.write_text(r'''
#include <stdio.h>
#include <string.h>
#include "cardioflow.h"
int main(void) {
    int h = 0, w = 0;
    if (cf_version() < 100) return 1;
    if (cf_pyramid_level_dims(32, 32, 2, &h, &w) != CF_OK || h != 8 || w != 8) return 2;
    if (cf_corr_pyramid_bytes(29, 32, 32, 4) != (size_t)161546240) return 3;
    if (cf_warp2d(NULL, NULL, 1, 1, 8, 8, CF_MODE_BILINEAR, NULL, NULL) != CF_ERR_BAD_ARG) return 4;
    if (strstr(cf_last_error(), "null") == NULL) return 5;
    if (cf_upsample_convex((const float*)256, (const float*)512, 1, 2, 8, 8, 4, 4.0f, (float*)1024, NULL) != CF_ERR_UNSUPPORTED) return 6;
    printf("c host ok\n");
    return 0;
}
''')
    exe = tmp_path / "host"
    cmd = ["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", os.path.join(root, "include"), str(src), "-o", str(exe),
           "-L", lib_dir, "-lcardioflow", "-Wl,-rpath," + lib_dir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0 and "c host ok" in r.stdout, (r.returncode, r.stdout, r.stderr)


def test_ncc_host_contract_without_a_gpu():
    """cardioflow_b200.NCC keeps the constructor / input contract of nnunet/lib/loss.py:15-76 and has no CPU path."""
    a = torch.rand(2, 1, 16, 16)
    loss = cf.NCC()
    assert (loss.win, loss.eps, loss.reduction) == (9, 1e-3, "mean") and len(list(loss.parameters())) == 0
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        loss(a, a)
    with pytest.raises(ValueError, match="5-D"):
        cf.NCC(reduction=None)(a, a)
    with pytest.raises(ValueError, match="single-channel"):
        loss(torch.rand(2, 3, 16, 16), torch.rand(2, 3, 16, 16))
    with pytest.raises(ValueError, match="differ"):
        loss(a, torch.rand(2, 1, 16, 17))
    lib = _cabi.lib()
    assert lib.cf_ncc(None, None, 1, 8, 8, 9, 1e-3, None, None, None, 0, None) == _cabi.CF_ERR_BAD_ARG
    assert lib.cf_ncc(ctypes.c_void_p(256), ctypes.c_void_p(512), 1, 8, 8, 7, 1e-3, ctypes.c_void_p(1024), None, None, 0, None) \
        == _cabi.CF_ERR_UNSUPPORTED and b"win 7" in lib.cf_last_error()
    assert lib.cf_ncc_workspace_bytes(300, 224, 224, 0) % 256 == 0 and lib.cf_ncc_workspace_bytes(0, 8, 8, 1) == 0
    assert lib.cf_sum_f64(None, 4, None, None, 0, None) == _cabi.CF_ERR_BAD_ARG and lib.cf_sum_f64_workspace_bytes() >= 8


def test_round2_entry_points_validate_their_arguments_without_a_gpu():
    """Null pointers / impossible shapes are rejected by the C ABI itself (CF_ERR_BAD_ARG or CF_ERR_UNSUPPORTED, with a message that
    names the call) before anything touches CUDA: the checks run here, on a box without a GPU."""
    import ctypes
    lib = _cabi.lib()
    N = None
    used, tiled = ctypes.c_int(0), ctypes.c_uint(0)
    one = ctypes.c_void_p(256)          # a non-null pointer that is never dereferenced: every call below fails validation first
    calls = {
        "cf_corr_build_ex": lambda: lib.cf_corr_build_ex(N, N, 0, 0, 1, 8, 8, 8, 2, 0, 0, N, N, 0, ctypes.byref(used), ctypes.byref(tiled), N),
        "cf_corr_build_ex/shape": lambda: lib.cf_corr_build_ex(one, one, 0, 0, 0, 8, 8, 8, 2, 0, 0, one, N, 0, ctypes.byref(used), ctypes.byref(tiled), N),
        "cf_corr_volume_bwd_ex": lambda: lib.cf_corr_volume_bwd_ex(N, N, N, 1, 8, 8, 8, 0, N, N, N, 0, ctypes.byref(used), N),
        "cf_corr_volume_bwd_ex/no output": lambda: lib.cf_corr_volume_bwd_ex(one, one, one, 1, 8, 8, 8, 0, N, N, N, 0, ctypes.byref(used), N),
        "cf_corr_volume_bwd_ex/precision": lambda: lib.cf_corr_volume_bwd_ex(one, one, one, 1, 8, 8, 8, _cabi.PREC_BF16, one, N, N, 0, ctypes.byref(used), N),   # plain bf16 is not a backward precision
        "cf_corr_volume_bwd_noalloc": lambda: lib.cf_corr_volume_bwd_noalloc(N, N, N, 1, 8, 8, 8, N, N, N),
        "cf_corr_volume_bwd_noalloc/shape": lambda: lib.cf_corr_volume_bwd_noalloc(one, one, one, 1, 0, 8, 8, one, N, N),
        "cf_corr_volume_bwd": lambda: lib.cf_corr_volume_bwd(N, N, N, 1, 8, 8, 8, N, N, N),
    }
    for name, call in calls.items():
        rc = call()
        msg = lib.cf_last_error().decode()
        assert rc in (_cabi.CF_ERR_BAD_ARG, _cabi.CF_ERR_UNSUPPORTED), (name, rc, msg)
        assert msg, name
        if "/" not in name:             # null-pointer checks name the call (cf_corr_build_ex reports as the cf_corr_build family)
            assert name.replace("_ex", "") in msg, (name, msg)
    # size queries answer without a GPU as well
    assert lib.cf_corr_volume_bwd_workspace_bytes(1, 8, 5, 5) == 0          # H*W = 25: no tensor-core path
    assert lib.cf_corr_volume_bwd_workspace_bytes(0, 8, 8, 8) == 0
