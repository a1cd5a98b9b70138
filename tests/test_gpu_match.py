"""GPU parity of the matching pipeline around the two kernels (SURVEY 8(f) f2,
mast3r_slam/matching.py:8-110): frodo_match_prep / _occlusion / frodo_convert_f32_f16 /
_finish and the single-call frodo_match, through the C ABI (frodo_slam_b200.matching).

Bars: every stage bit-identical to the oracle (oracle/oracle.py fixes the operation order);
against the reference's own Python (tests/golden/golden_match.npz) the float prologue agrees
to 2 ulp (torch sums in another order) and the final indices and masks are identical."""
import os

import numpy as np
import pytest
import torch

from oracle import oracle
from frodo_slam_b200 import synth

pytestmark = pytest.mark.gpu

GOLD_MATCH = os.path.join(os.path.dirname(__file__), "golden", "golden_match.npz")
CFG = dict(synth.MATCH_PARAMS)


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def fm():
    from frodo_slam_b200 import matching
    return matching


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _inputs(B, H, W, seed, init):
    d = synth.matching_pairs(B, H, W, seed=seed, device="cpu")
    X11, X21 = d["X11"].float().numpy(), d["X21"].float().numpy()
    D11, D21 = d["D11"].float().numpy(), d["D21"].float().numpy()
    idx0 = None
    if init == "jitter":
        g = torch.Generator().manual_seed(seed + 9)
        ident = torch.arange(H * W)[None].repeat(B, 1)
        idx0 = (ident + torch.randint(-3, 4, (B, H * W), generator=g)
                + W * torch.randint(-3, 4, (B, H * W), generator=g)).clamp(0, H * W - 1).numpy()
    elif init == "random":
        g = torch.Generator().manual_seed(seed + 9)
        idx0 = torch.randint(0, H * W, (B, H * W), generator=g).numpy()
    return X11, X21, D11, D21, idx0


# ------------------------------------------------------------------ golden replay
@pytest.mark.parametrize("tag", ["a", "b"])
def test_golden_match_replayed_on_gpu(fm, dev, tag):
    g = np.load(GOLD_MATCH)
    cfg = {k[4:]: float(g[k]) for k in g.files if k.startswith("cfg_")}
    init = None if tag == "a" else _t(g["idx_init"], dev)
    X11, X21 = _t(g["X11"], dev), _t(g["X21"], dev)
    D11, D21 = _t(g["D11"].astype(np.float32), dev), _t(g["D21"].astype(np.float32), dev)
    rays, pts, p0 = fm.prep_for_iter_proj(X11, X21, init)
    assert (rays.cpu().numpy() - g[tag + "_rays_with_grad_img"]).__abs__().max() <= 2.4e-7
    assert (pts.cpu().numpy() - g[tag + "_pts3d_norm"]).__abs__().max() <= 2.4e-7
    assert np.array_equal(p0.cpu().numpy(), g[tag + "_p_init"])
    idx, valid = fm.match(X11, X21, D11, D21, init, cfg=cfg)
    assert idx.dtype == torch.int64 and valid.dtype == torch.bool and valid.shape[-1] == 1
    assert np.array_equal(idx.cpu().numpy(), g[tag + "_idx_1_to_2"])
    assert np.array_equal(valid.cpu().numpy(), g[tag + "_valid"])


# ------------------------------------------------------------------ stages vs the oracle
@pytest.mark.parametrize("B,H,W,init", [(2, 32, 48, None), (1, 37, 53, "jitter"),
                                        (3, 16, 130, "random"), (1, 9, 33, None)])
def test_prep_bit_exact_vs_oracle(fm, dev, B, H, W, init):
    X11, X21, _, _, idx0 = _inputs(B, H, W, 11, init)
    r_o, p_o, i_o = oracle.prep_for_iter_proj(X11, X21, idx0)
    rays, pts, p0 = fm.prep_for_iter_proj(_t(X11, dev), _t(X21, dev),
                                          None if idx0 is None else _t(idx0, dev))
    assert np.array_equal(rays.cpu().numpy(), r_o)
    assert np.array_equal(pts.cpu().numpy(), p_o)
    assert np.array_equal(p0.cpu().numpy(), i_o)


def test_prep_zero_vectors_do_not_divide_by_zero(fm, dev):
    X11, X21, _, _, _ = _inputs(1, 16, 32, 2, None)
    X11[0, 3:5, 7:9] = 0.0   # F.normalize: x / max(|x|, 1e-12) -> 0
    X21[0, 0, 0] = 0.0
    r_o, p_o, _ = oracle.prep_for_iter_proj(X11, X21, None)
    rays, pts, _ = fm.prep_for_iter_proj(_t(X11, dev), _t(X21, dev))
    assert np.isfinite(rays.cpu().numpy()).all()
    assert np.array_equal(rays.cpu().numpy(), r_o) and np.array_equal(pts.cpu().numpy(), p_o)


@pytest.mark.parametrize("n", [0, 1, 7, 8, 4099, 24 * 48 * 64])
def test_convert_equals_torch_half(dev, n):
    import ctypes
    from frodo_slam_b200 import _lib
    g = torch.Generator().manual_seed(n)
    x = torch.randn(n + 3, generator=g) * 10.0
    if n >= 8:
        x[:8] = torch.tensor([float("inf"), -float("inf"), float("nan"), 65504.0, 65520.0, 1e-8,
                              6.0e-8, -0.0])
    x = x.to(dev)
    for off in (0, 1):   # aligned and unaligned start (scalar path)
        src = x[off:off + n]
        dst = torch.full((n + 1,), 7.0, dtype=torch.float16, device=dev)
        lib = _lib.load()
        _lib.check(lib.frodo_convert_f32_f16(
            ctypes.c_void_p(src.data_ptr()), ctypes.c_void_p(dst.data_ptr()), n,
            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
        want = src.half()
        assert torch.equal(dst[:n].view(torch.int16), want.view(torch.int16))
        assert float(dst[n]) == 7.0   # no write past the end


@pytest.mark.parametrize("B,H,W,init", [(2, 32, 48, None), (1, 37, 53, "jitter"),
                                        (2, 24, 40, "random")])
def test_match_vs_oracle_and_staged_equals_single_call(fm, dev, B, H, W, init):
    X11, X21, D11, D21, idx0 = _inputs(B, H, W, 21, init)
    idx_o, valid_o = oracle.match(X11, X21, D11, D21, idx0, **CFG)
    args = (_t(X11, dev), _t(X21, dev), _t(D11, dev), _t(D21, dev),
            None if idx0 is None else _t(idx0, dev))
    idx, valid = fm.match(*args, cfg=CFG)
    assert np.array_equal(idx.cpu().numpy(), idx_o)
    assert np.array_equal(valid.cpu().numpy(), valid_o)
    idx_s, valid_s = fm.match_iterative_proj(*args, cfg=CFG)
    assert torch.equal(idx_s, idx) and torch.equal(valid_s, valid)
    assert 0.05 < valid_o.mean() < 0.999


def test_match_full_size_vs_oracle(fm, dev):
    """BASELINE.json's 512x384, one pair: prologue bit-exact, final indices / masks identical."""
    H, W = 384, 512
    X11, X21, D11, D21, _ = _inputs(1, H, W, 5, None)
    r_o, p_o, _ = oracle.prep_for_iter_proj(X11, X21, None)
    rays, pts, _ = fm.prep_for_iter_proj(_t(X11, dev), _t(X21, dev))
    assert np.array_equal(rays.cpu().numpy(), r_o) and np.array_equal(pts.cpu().numpy(), p_o)
    idx_o, valid_o = oracle.match(X11, X21, D11, D21, None, **CFG)
    idx, valid = fm.match(_t(X11, dev), _t(X21, dev), _t(D11, dev), _t(D21, dev), cfg=CFG)
    assert np.array_equal(idx.cpu().numpy(), idx_o)
    assert np.array_equal(valid.cpu().numpy(), valid_o)
    # size-independent properties: indices inside the image, masked-in matches are close in 3-D
    i = idx.cpu().numpy()[0]
    assert i.min() >= 0 and i.max() < H * W
    v = valid.cpu().numpy()[0, :, 0]
    d = np.linalg.norm(X11.reshape(-1, 3)[i[v]] - X21.reshape(-1, 3)[v], axis=-1)
    assert v.any() and np.percentile(d, 99) < 0.5


def test_match_determinism_and_batch_independence(fm, dev):
    X11, X21, D11, D21, _ = _inputs(4, 32, 64, 8, None)
    a = [_t(x, dev) for x in (X11, X21, D11, D21)]
    i1, v1 = fm.match(*a, cfg=CFG)
    i2, v2 = fm.match(*a, cfg=CFG)
    assert torch.equal(i1, i2) and torch.equal(v1, v2)
    b = [x[2:3].contiguous() for x in a]
    i3, v3 = fm.match(*b, cfg=CFG)
    assert torch.equal(i3, i1[2:3]) and torch.equal(v3, v1[2:3])


def test_match_argument_errors(fm, dev):
    X11, X21, D11, D21, _ = _inputs(1, 16, 32, 1, None)
    a = [_t(x, dev) for x in (X11, X21, D11, D21)]
    with pytest.raises(RuntimeError):
        fm.match(a[0].cpu(), a[1], a[2], a[3])            # no CPU path
    with pytest.raises(RuntimeError):
        fm.match(a[0].transpose(1, 2), a[1], a[2], a[3])   # non-contiguous
    with pytest.raises(RuntimeError):
        fm.match(a[0].double(), a[1], a[2], a[3])
    idx, valid = fm.match(a[0][:0], a[1][:0], a[2][:0], a[3][:0])   # empty batch
    assert idx.shape == (0, 16 * 32) and valid.shape == (0, 16 * 32, 1)


# ------------------------------------------------ match_quality (tracker.py:194-206, global_opt.py:98-112)

def _quality_inputs(B, N, N_src, seed):
    g = torch.Generator().manual_seed(seed)
    Qs = torch.rand(B, N_src, 1, generator=g) * 3.0
    Qd = torch.rand(B, N, 1, generator=g) * 3.0
    Cs = torch.rand(B, N_src, 1, generator=g) * 2.0 - 0.3
    Cd = torch.rand(B, N, 1, generator=g) * 2.0 - 0.3
    idx = torch.randint(0, N_src, (B, N), generator=g)
    vm = torch.rand(B, N, 1, generator=g) < 0.8
    return Qs, Qd, Cs, Cd, idx, vm


@pytest.mark.parametrize("B,N,N_src", [(1, 48 * 64, 48 * 64), (3, 1001, 777), (2, 5, 9),
                                       (1, 384 * 512, 384 * 512)])
@pytest.mark.parametrize("with_conf", [True, False])
def test_match_quality_bitwise_vs_torch_and_oracle(fm, dev, B, N, N_src, with_conf):
    Qs, Qd, Cs, Cd, idx, vm = _quality_inputs(B, N, N_src, seed=B * 7 + N)
    Q_conf, C_conf = 1.5, 0.25
    # the reference's expressions (tracker.py:194-206), evaluated by torch ON THE GPU as the
    # reference does (torch's CPU sqrt is not correctly rounded, CUDA's is)
    rows = torch.arange(B, device=dev)[:, None]
    Qs_d, Qd_d, Cs_d, Cd_d, idx_d, vm_d = (x.to(dev) for x in (Qs, Qd, Cs, Cd, idx, vm))
    Q_t = torch.sqrt(Qs_d[rows, idx_d] * Qd_d)
    vq_t = vm_d & (Q_t > Q_conf)
    vo_t = vq_t & (Cs_d[rows, idx_d] > C_conf) & (Cd_d > C_conf) if with_conf else vq_t
    Q_t, vq_t, vo_t = Q_t.cpu(), vq_t.cpu(), vo_t.cpu()
    Q, vo, vq, counts = fm.match_quality(Qs.to(dev), Qd.to(dev), idx.to(dev), vm.to(dev), Q_conf,
                                         Cs.to(dev) if with_conf else None,
                                         Cd.to(dev) if with_conf else None, C_conf)
    assert Q.shape == Qd.shape and vo.shape == vm.shape
    assert torch.equal(Q.cpu(), Q_t) and torch.equal(vq.cpu(), vq_t) and torch.equal(vo.cpu(), vo_t)
    assert counts.cpu().tolist() == [[int(vo_t[b].sum()), int(vq_t[b].sum())] for b in range(B)]
    o = oracle.match_quality(Qs.numpy(), Qd.numpy(), idx.numpy(), vm.numpy(), Q_conf,
                             Cs.numpy() if with_conf else None, Cd.numpy() if with_conf else None,
                             C_conf)
    assert np.array_equal(Q.cpu().numpy().reshape(B, N), o[0])
    assert np.array_equal(vo.cpu().numpy().reshape(B, N), o[1])
    assert np.array_equal(counts.cpu().numpy(), o[3])


def test_match_quality_errors_and_empty(fm, dev):
    Qs, Qd, Cs, Cd, idx, vm = _quality_inputs(2, 16, 16, seed=1)
    with pytest.raises(RuntimeError):  # CPU tensors
        fm.match_quality(Qs, Qd, idx, vm, 1.5)
    with pytest.raises(RuntimeError):  # wrong size
        fm.match_quality(Qs.to(dev), Qd.to(dev)[:, :8].contiguous(), idx.to(dev), vm.to(dev), 1.5)
    e = fm.match_quality(Qs.to(dev)[:0], Qd.to(dev)[:0], idx.to(dev)[:0], vm.to(dev)[:0], 1.5)
    assert e[0].numel() == 0 and e[3].shape == (0, 2)
