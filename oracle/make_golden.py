"""Generate the golden fixtures under tests/golden/ from the REAL reference code (test infrastructure).

Run in the build container only (``/root/reference`` does not exist on the GPU box):

    python oracle/make_golden.py

* warp fixtures:  outputs of the reference's own ``SpatialTransformer`` / ``VecInt`` /
  ``SpatialTransformerContour``, imported by file path from
  ``/root/reference/nnunet/network_architecture/integration.py`` (never ``import nnunet``: its
  ``__init__`` pulls packages that are absent here).
* corr fixtures:  outputs of torchvision's ``CorrBlock`` (``models/optical_flow/raft.py``), the in-tree
  third-party call site of the reference (``nnMTLTrainerV2Raft.py:97-118``); the reference's own
  ``nnunet/lib/raft_initial.py`` was never committed.

* upsample fixtures:  outputs of the reference's own ``upsample_flow`` / ``upsample_seg`` methods, compiled out of
  ``/root/reference/nnunet/network_architecture/SegFlowGaussian.py:846-871`` by name (the module's imports are absent).

Every fixture stores its inputs next to the outputs, so the tests do not depend on RNG reproducibility.
"""
import importlib.util
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(os.path.dirname(HERE), "tests", "golden")
REF_FILE = "/root/reference/nnunet/network_architecture/integration.py"


def load_reference_integration():
    spec = importlib.util.spec_from_file_location("ref_integration", REF_FILE)
    mod = importlib.util.module_from_spec(spec)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        spec.loader.exec_module(mod)
    return mod


def gen(seed):
    g = torch.Generator()
    g.manual_seed(seed)
    return g


def warp_cases():
    """name -> (size, B, C, flow builder)."""
    def smooth(g, b, h, w, amp):
        coarse = torch.randn(b, 2, 5, 5, generator=g) * amp
        return torch.nn.functional.interpolate(coarse, size=(h, w), mode="bicubic", align_corners=True)

    cases = {}
    g = gen(2001)
    cases["zero_24"] = ((24, 24), torch.zeros(2, 2, 24, 24), torch.rand(2, 3, 24, 24, generator=g))
    cases["integer_32x40"] = ((32, 40), torch.randint(-3, 4, (2, 2, 32, 40), generator=g).float(),
                              torch.rand(2, 1, 32, 40, generator=g))
    cases["half_16"] = ((16, 16), torch.full((1, 2, 16, 16), 0.5), torch.rand(1, 4, 16, 16, generator=g))
    cases["noise_48"] = ((48, 48), torch.randn(2, 2, 48, 48, generator=g) * 3.0, torch.rand(2, 5, 48, 48, generator=g))
    cases["smooth_64x48"] = ((64, 48), smooth(g, 2, 64, 48, 4.0), torch.rand(2, 2, 64, 48, generator=g))
    cases["oob_20"] = ((20, 20), torch.randn(1, 2, 20, 20, generator=g) * 30.0, torch.rand(1, 2, 20, 20, generator=g))
    return cases


def make_warp(ref):
    out = {}
    for name, (size, flow, src) in warp_cases().items():
        st = ref.SpatialTransformer(size)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            res = st(flow=flow.clone(), original=src.clone())
            near = ref.SpatialTransformer(size, mode="nearest")(flow.clone(), src.clone())
        out[f"{name}.flow"] = flow.numpy()
        out[f"{name}.src"] = src.numpy()
        out[f"{name}.bilinear"] = res.numpy()
        out[f"{name}.nearest"] = near.numpy()
    # state-dict contract
    st = ref.SpatialTransformer((6, 5))
    out["state.grid_6x5"] = st.state_dict()["grid"].numpy()
    # one-hot + warp + argmax pattern (SegFlowGaussian.py:3571-3580)
    g = gen(2002)
    label = torch.randint(0, 4, (2, 1, 40, 40), generator=g)
    flow = torch.randn(2, 2, 40, 40, generator=g) * 2.0
    onehot = torch.nn.functional.one_hot(label[:, 0].long(), num_classes=4).permute(0, 3, 1, 2).contiguous().float()
    reg = ref.SpatialTransformer((40, 40))(flow=flow.clone(), original=onehot, mode="bilinear")
    out["argmax.label"] = label.numpy()
    out["argmax.flow"] = flow.numpy()
    out["argmax.out"] = torch.argmax(reg, dim=1, keepdim=True).numpy()
    # VecInt (integration.py:82-99)
    vec = torch.randn(1, 2, 32, 32, generator=g)
    out["vecint.vec"] = vec.numpy()
    out["vecint.out7"] = ref.VecInt((32, 32), 7)(vec.clone()).numpy()
    # SpatialTransformerContour (integration.py:5-34): P points sampled from a flow field
    locs = torch.rand(1, 2, 1, 50, generator=g) * 31.0
    field = torch.randn(1, 2, 32, 32, generator=g)
    out["contour.locs"] = locs.numpy()
    out["contour.field"] = field.numpy()
    out["contour.out"] = ref.SpatialTransformerContour((32, 32))(locs.clone(), field).numpy()
    np.savez_compressed(os.path.join(GOLDEN, "warp_reference.npz"), **out)
    print("warp_reference.npz:", len(out), "arrays")


def coords_trajectory(g, b, h, w, iters, sigma=0.5):
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    grid = torch.stack([xs, ys]).float()[None].repeat(b, 1, 1, 1)
    out, cur = [], grid.clone()
    for k in range(iters):
        out.append(cur.clone())
        cur = cur + torch.randn(b, 2, h, w, generator=g) * sigma * (k + 1)
        cur[:, 0].clamp_(-4, w + 4)
        cur[:, 1].clamp_(-4, h + 4)
    return out


def make_corr():
    from torchvision.models.optical_flow.raft import CorrBlock as TVCorrBlock
    out = {}
    for name, (b, c, h, w, seed) in {"c16_16x16": (1, 16, 16, 16, 3001),
                                     "c8_24x24": (1, 8, 24, 24, 3002),
                                     "c32_16x24": (2, 32, 16, 24, 3003)}.items():
        g = gen(seed)
        f1 = torch.randn(b, c, h, w, generator=g)
        f2 = torch.randn(b, c, h, w, generator=g)
        blk = TVCorrBlock(num_levels=4, radius=4)
        blk.build_pyramid(f1, f2)
        traj = coords_trajectory(g, b, h, w, 3)
        traj.append(traj[0] + 1000.0)  # far out of range -> all zeros
        out[f"{name}.f1"] = f1.numpy()
        out[f"{name}.f2"] = f2.numpy()
        for lvl in range(1, 4):
            out[f"{name}.pyr{lvl}"] = blk.corr_pyramid[lvl].numpy()
        for k, coords in enumerate(traj):
            out[f"{name}.coords{k}"] = coords.numpy()
            out[f"{name}.lookup{k}"] = blk.index_pyramid(coords).numpy()
    np.savez_compressed(os.path.join(GOLDEN, "corr_torchvision.npz"), **out)
    print("corr_torchvision.npz:", len(out), "arrays")


def load_reference_methods(path, names):
    """Compile single methods of a reference class out of its source file (the module itself cannot be imported: its
    imports need packages that are absent here).  The methods used here touch only torch / F."""
    import ast
    import torch.nn.functional as F
    src = open(path).read()
    tree = ast.parse(src)
    found = {}
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name in names and node.name not in found:
            ns = {"torch": torch, "F": F}
            exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
            found[node.name] = ns[node.name]
    missing = set(names) - set(found)
    if missing:
        raise RuntimeError(f"{path}: methods not found: {sorted(missing)}")
    return found


def make_utils_warp():
    """lib/utils.py:29-59 `warp`, compiled out of the reference file; its hard-coded `grid.cuda()` is made a no-op for
    the duration of the call (this container has no GPU; on CPU the function is otherwise unchanged)."""
    import torch.nn as nn
    import ast
    path = "/root/reference/nnunet/lib/utils.py"
    tree = ast.parse(open(path).read())
    node = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "warp")
    ns = {"torch": torch, "nn": nn}
    exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
    out = {}
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        for name, (b, c, h, w, amp, seed) in {"uw_24": (2, 3, 24, 24, 2.0, 5001), "uw_17x40": (1, 2, 17, 40, 4.0, 5002),
                                              "uw_oob_16": (1, 1, 16, 16, 25.0, 5003), "uw_zero_12": (1, 2, 12, 12, 0.0, 5004)}.items():
            g = gen(seed)
            x = torch.rand(b, c, h, w, generator=g)
            flo = torch.randn(b, 2, h, w, generator=g) * amp
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                res = ns["warp"](x.clone(), flo.clone())
            out[f"{name}.x"], out[f"{name}.flo"], out[f"{name}.out"] = x.numpy(), flo.numpy(), res.numpy()
    finally:
        torch.Tensor.cuda = orig_cuda
    np.savez_compressed(os.path.join(GOLDEN, "utils_warp_reference.npz"), **out)
    print("utils_warp_reference.npz:", len(out), "arrays")


def make_ncc():
    """lib/loss.py:15-76 `class NCC`, compiled out of the reference file (the module imports kornia / matplotlib, absent
    here); its hard-coded `win_size.cuda()` is a no-op for the duration of the calls.  Cases: noise, smooth (ill-conditioned
    variances), a warped pair, borders smaller than the window, and the 5-D reduction=None branch."""
    import ast
    import torch.nn.functional as F
    path = "/root/reference/nnunet/lib/loss.py"
    tree = ast.parse(open(path).read())
    node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "NCC")
    ns = {"torch": torch, "F": F, "np": np}
    exec(compile(ast.Module(body=[node], type_ignores=[]), path, "exec"), ns)
    NCC = ns["NCC"]
    out = {}
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        def smooth(g, n, h, w):
            z = torch.randn(n, 1, max(h // 8, 2), max(w // 8, 2), generator=g)
            return F.interpolate(z, size=(h, w), mode="bicubic", align_corners=True) * 0.3 + 0.5
        cases = {}
        g = gen(6001); cases["ncc_noise_40"] = (torch.rand(2, 1, 40, 40, generator=g), torch.rand(2, 1, 40, 40, generator=g))
        g = gen(6002); a = smooth(g, 2, 48, 37); cases["ncc_smooth_48x37"] = (a, a + 0.05 * torch.randn(2, 1, 48, 37, generator=g))
        g = gen(6003); a = smooth(g, 1, 64, 64); cases["ncc_shift_64"] = (a, torch.roll(a, shifts=(2, -3), dims=(2, 3)))
        g = gen(6004); cases["ncc_tiny_5x7"] = (torch.rand(3, 1, 5, 7, generator=g), torch.rand(3, 1, 5, 7, generator=g))
        g = gen(6005); a = torch.rand(1, 1, 33, 70, generator=g); cases["ncc_same_33x70"] = (a, a.clone())
        for name, (I, J) in cases.items():
            out[f"{name}.I"], out[f"{name}.J"] = I.numpy(), J.numpy()
            out[f"{name}.loss"] = NCC()(I.clone(), J.clone()).numpy()
            T = I.shape[0]
            out[f"{name}.map"] = NCC(reduction=None)(I.clone()[:, None], J.clone()[:, None]).numpy()     # [T,1,1,H,W]
        # gradients of the mean loss w.r.t. both images (autograd through the reference's own expression)
        for name in ("ncc_noise_40", "ncc_smooth_48x37"):
            I = torch.from_numpy(out[f"{name}.I"]).clone().requires_grad_(True)
            J = torch.from_numpy(out[f"{name}.J"]).clone().requires_grad_(True)
            NCC()(I, J).backward()
            out[f"{name}.gI"], out[f"{name}.gJ"] = I.grad.numpy(), J.grad.numpy()
    finally:
        torch.Tensor.cuda = orig_cuda
    np.savez_compressed(os.path.join(GOLDEN, "ncc_reference.npz"), **out)
    print("ncc_reference.npz:", len(out), "arrays")


def make_upsample():
    ref = load_reference_methods("/root/reference/nnunet/network_architecture/SegFlowGaussian.py",
                                 ["upsample_flow", "upsample_seg"])
    out = {}
    for name, (n, h, w, seed) in {"u_6x5": (2, 6, 5, 4001), "u_12x12": (1, 12, 12, 4002), "u_9x33": (1, 9, 33, 4003)}.items():
        g = gen(seed)
        flow = torch.randn(n, 2, h, w, generator=g) * 2.0
        seg = torch.randn(n, 4, h, w, generator=g)
        mask = torch.randn(n, 576, h, w, generator=g) * 3.0
        out[f"{name}.flow"], out[f"{name}.seg"], out[f"{name}.mask"] = flow.numpy(), seg.numpy(), mask.numpy()
        out[f"{name}.up_flow"] = ref["upsample_flow"](None, flow.clone(), mask.clone()).numpy()
        out[f"{name}.up_seg"] = ref["upsample_seg"](None, seg.clone(), mask.clone()).numpy()
    np.savez_compressed(os.path.join(GOLDEN, "upsample_reference.npz"), **out)
    print("upsample_reference.npz:", len(out), "arrays")


if __name__ == "__main__":
    if not os.path.exists(REF_FILE):
        sys.exit("reference not present: fixtures can only be regenerated in the build container")
    os.makedirs(GOLDEN, exist_ok=True)
    torch.set_num_threads(1)
    make_warp(load_reference_integration())
    make_corr()
    make_upsample()
    make_utils_warp()
    make_ncc()
