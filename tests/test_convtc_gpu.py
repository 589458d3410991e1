"""tcgen05 / TMA / TMEM implicit-GEMM convolution (bf16) through the plan API of the C ABI, against
the CPU oracle on the same bf16-rounded inputs.  Tolerance: the north star's 2e-2 for bf16 mode
(measured error is the bf16 output rounding, ~4e-3)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

from util import rel_err, dev, stream, lib, check, ptr, ConvDesc, RN_BF16  # noqa: E402
from oracle import resnet_oracle as ro  # noqa: E402

BF = torch.bfloat16

CASES = [
    # N, H, W, Cin, Cout, k, stride, pad
    (4, 8, 8, 64, 64, 1, 1, 0),        # plain GEMM, one k-chunk
    (4, 8, 8, 64, 256, 1, 1, 0),       # N = 256 tile
    (2, 16, 16, 256, 64, 1, 1, 0),     # 4 k-chunks
    (4, 16, 16, 64, 64, 3, 1, 1),      # 9 taps, halo zero fill
    (8, 14, 14, 256, 256, 3, 1, 1),    # C3 stage-3 shape
    (6, 7, 7, 512, 512, 3, 1, 1),      # C3 stage-4 shape: box 7x7x2 (98 of 128 rows), ragged N
    (2, 56, 56, 64, 64, 3, 1, 1),      # C3 stage-1 shape
    (4, 32, 32, 16, 16, 3, 1, 1),      # CIFAR: 32-byte rows (SW32)
    (4, 16, 16, 32, 32, 3, 1, 1),      # CIFAR: 64-byte rows (SW64)
    (3, 9, 11, 24, 48, 3, 1, 1),       # ragged everything: K tail zero-filled by TMA
    (2, 32, 32, 160, 160, 3, 1, 1),    # WRN width: N = 160 tile, 32-channel store groups, K tail
    (2, 8, 8, 2048, 512, 1, 1, 0),     # deep K
]


def _weights(w):
    k, _, cin, cout = w.shape
    wf = torch.empty(k * k * cout * cin, dtype=BF, device="cuda")
    wd = torch.empty(k * k * cin * cout, dtype=BF, device="cuda")
    W32 = dev(w.float())
    check(lib.rn_weights_to_bf16(ptr(W32), k, cin, cout, ptr(wf), ptr(wd), stream()), "weights_to_bf16")
    torch.cuda.synchronize()
    assert torch.equal(wd.view(k, k, cin, cout).cpu(), w.to(BF))
    assert torch.equal(wf.view(k, k, cout, cin).cpu(), w.to(BF).permute(0, 1, 3, 2))
    return wf, wd


@pytest.mark.parametrize("case", CASES)
def test_convtc_fprop_and_dgrad(case):
    N, H, W, Cin, Cout, k, s, pad = case
    Ho, Wo = H, W
    g = torch.Generator().manual_seed(sum(case))
    x = torch.randn(N, H, W, Cin, generator=g).to(BF)
    w = (torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5).to(BF)
    dy = torch.randn(N, Ho, Wo, Cout, generator=g).to(BF)
    net = ro.Net(dtype=torch.float64, params={"w": w.double().numpy()})
    xv = ro.Var(x.double())
    yv = net.conv2d(xv, "w", k, Cin, Cout, s, pad, (Ho, Wo), None, "y")
    yv.grad = dy.double()
    net.backward()
    d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, k, s, pad, RN_BF16)
    assert lib.rn_convtc_supported(C.byref(d), 0) == 1
    kinds = (0, 1) if lib.rn_convtc_supported(C.byref(d), 1) else (0,)    # 24 output columns cannot form a UMMA N tile
    assert len(kinds) == 2 or Cin % 16
    wf, wd = _weights(w)
    X, DY = dev(x), dev(dy)
    Y = torch.full((N, Ho, Wo, Cout), float("nan"), dtype=BF, device="cuda")
    DX = torch.full((N, H, W, Cin), float("nan"), dtype=BF, device="cuda")
    for kind, a, b, out, want in ((0, X, wd, Y, yv.val), (1, DY, wd, DX, xv.grad)):
        if kind not in kinds:
            continue
        plan = C.c_void_p()
        check(lib.rn_convtc_plan_create(C.byref(d), kind, ptr(a), ptr(b), ptr(out), None, 0, C.byref(plan)), "plan")
        for _ in range(2):      # run twice: barriers / TMEM are per-launch state
            check(lib.rn_convtc_run(plan, 0.0, stream()), "run")
        torch.cuda.synchronize()
        lib.rn_convtc_plan_destroy(plan)
        assert torch.isfinite(out.float()).all(), "unwritten outputs (kind %d)" % kind
        assert rel_err(out.float(), want) < 2e-2, ("kind", kind)


PAIR_CASES = [CASES[1], CASES[4], CASES[5], CASES[11], (8, 14, 14, 1024, 256, 1, 1, 0), (5, 14, 14, 256, 512, 3, 1, 1)]


@pytest.mark.parametrize("mode", ["1", "0"])
@pytest.mark.parametrize("case", PAIR_CASES)
def test_convtc_cta_pairs(case, mode, monkeypatch):
    """tcgen05.mma.cta_group::2: clusters of two CTAs, M = 256 per instruction, each CTA stages half of the weight
    tile.  RN_TC_PAIR=1 forces pairs wherever the tile allows, =0 forbids them; both must match the oracle, including
    an odd number of pixel-tile groups (the last pair's second CTA works on an out-of-range tile) and N = 512
    (two column tiles per pair)."""
    from distributed_tensorflow_resnet_b200._lib import describe_plan
    monkeypatch.setenv("RN_TC_PAIR", mode)
    N, H, W, Cin, Cout, k, s, pad = case
    g = torch.Generator().manual_seed(sum(case) + 7)
    x = torch.randn(N, H, W, Cin, generator=g).to(BF)
    w = (torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5).to(BF)
    dy = torch.randn(N, H, W, Cout, generator=g).to(BF)
    net = ro.Net(dtype=torch.float64, params={"w": w.double().numpy()})
    xv = ro.Var(x.double())
    yv = net.conv2d(xv, "w", k, Cin, Cout, s, pad, (H, W), None, "y")
    yv.grad = dy.double()
    net.backward()
    d = ConvDesc(N, H, W, Cin, H, W, Cout, k, s, pad, RN_BF16)
    _, wd = _weights(w)
    X, DY = dev(x), dev(dy)
    Y = torch.full((N, H, W, Cout), float("nan"), dtype=BF, device="cuda")
    DX = torch.full((N, H, W, Cin), float("nan"), dtype=BF, device="cuda")
    for kind, a, out, want in ((0, X, Y, yv.val), (1, DY, DX, xv.grad)):
        plan = C.c_void_p()
        check(lib.rn_convtc_plan_create(C.byref(d), kind, ptr(a), ptr(wd), ptr(out), None, 0, C.byref(plan)), "plan")
        ncols = Cout if kind == 0 else Cin
        assert ("pair=1" in describe_plan(plan)) == (mode == "1" and ncols % 128 == 0), describe_plan(plan)
        for _ in range(2):
            check(lib.rn_convtc_run(plan, 0.0, stream()), "run")
        torch.cuda.synchronize()
        lib.rn_convtc_plan_destroy(plan)
        assert torch.isfinite(out.float()).all()
        assert rel_err(out.float(), want) < 2e-2, ("kind", kind)


HALO_CASES = [(2, 16, 16, 64, 64, 3, 1, 1), (2, 24, 24, 64, 64, 3, 1, 1), (3, 14, 14, 64, 128, 3, 1, 1),
              (2, 28, 28, 128, 128, 3, 1, 1), (2, 56, 56, 64, 64, 3, 1, 1), (5, 9, 11, 64, 64, 3, 1, 1)]


@pytest.mark.parametrize("pair", ["0", "1"])
@pytest.mark.parametrize("case", HALO_CASES)
def test_convtc_halo_mode(case, pair, monkeypatch):
    """3x3 halo mode (opt-in, RN_TC_HALO=1): one (BH + 2)-row box per horizontal tap, the three vertical taps read it
    at row offsets 0, BW, 2*BW through the matrix descriptor's start address — including box widths that are not a
    multiple of the 8-row swizzle atom (5, 7, 4, 14, 9 ...).  fprop and dgrad against the oracle, with and without
    CTA pairs."""
    from distributed_tensorflow_resnet_b200._lib import describe_plan
    monkeypatch.setenv("RN_TC_HALO", "1")
    monkeypatch.setenv("RN_TC_PAIR", pair)
    N, H, W, Cin, Cout, k, s, pad = case
    g = torch.Generator().manual_seed(sum(case) + 3)
    x = torch.randn(N, H, W, Cin, generator=g).to(BF)
    w = (torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5).to(BF)
    dy = torch.randn(N, H, W, Cout, generator=g).to(BF)
    net = ro.Net(dtype=torch.float64, params={"w": w.double().numpy()})
    xv = ro.Var(x.double())
    yv = net.conv2d(xv, "w", k, Cin, Cout, s, pad, (H, W), None, "y")
    yv.grad = dy.double()
    net.backward()
    d = ConvDesc(N, H, W, Cin, H, W, Cout, k, s, pad, RN_BF16)
    _, wd = _weights(w)
    X, DY = dev(x), dev(dy)
    Y = torch.full((N, H, W, Cout), float("nan"), dtype=BF, device="cuda")
    DX = torch.full((N, H, W, Cin), float("nan"), dtype=BF, device="cuda")
    for kind, a, out, want in ((0, X, Y, yv.val), (1, DY, DX, xv.grad)):
        plan = C.c_void_p()
        check(lib.rn_convtc_plan_create(C.byref(d), kind, ptr(a), ptr(wd), ptr(out), None, 0, C.byref(plan)), "plan")
        assert "halo=1" in describe_plan(plan), describe_plan(plan)
        for _ in range(2):
            check(lib.rn_convtc_run(plan, 0.0, stream()), "run")
        torch.cuda.synchronize()
        lib.rn_convtc_plan_destroy(plan)
        assert torch.isfinite(out.float()).all()
        assert rel_err(out.float(), want) < 2e-2, ("kind", kind)


S2_CASES = [
    # N, H, W, Cin, Cout, k, stride, pad, (Ho, Wo)
    (4, 16, 16, 64, 128, 3, 2, 1, (8, 8)),       # official 3x3 s2: symmetric fixed_padding + VALID
    (4, 16, 16, 64, 128, 3, 2, 0, (8, 8)),       # upstream 3x3 s2: TF-SAME (0 before, 1 after)
    (4, 16, 16, 64, 128, 1, 2, 0, (8, 8)),       # 1x1 s2 projection shortcut
    (2, 56, 56, 256, 512, 1, 2, 0, (28, 28)),    # C3 stage-2 projection
    (2, 28, 28, 128, 128, 3, 2, 1, (14, 14)),    # C3 3x3 s2 (stride on the 3x3 of the bottleneck)
    (3, 15, 13, 64, 64, 3, 2, 1, (8, 7)),        # odd sizes: parity views of different extents
    (8, 32, 32, 16, 32, 3, 2, 1, (16, 16)),      # CIFAR stage transition
    (8, 32, 32, 16, 32, 1, 2, 0, (16, 16)),
]

WGRAD_CASES = CASES[0:3] + CASES[3:7] + [CASES[11]] + [
    (16, 14, 14, 128, 128, 3, 1, 1),             # BLOCK_N = 128, four accumulators
    (4, 28, 28, 192, 64, 1, 1, 0),               # three 64-channel row blocks
    (8, 32, 32, 16, 16, 3, 1, 1),                # CIFAR widths: 16-channel boxes (32-byte MN-major rows), 8 taps per M tile
    (8, 16, 16, 32, 32, 3, 1, 1),                # 32-channel boxes (64-byte rows)
    (8, 16, 16, 16, 32, 1, 1, 0),                # mixed box widths A=16, B=32
    (4, 32, 32, 160, 160, 3, 1, 1),              # WRN-28-10 width: 32-channel boxes, N = 160
    (4, 16, 16, 32, 64, 3, 1, 1),
    (4, 16, 16, 160, 320, 3, 1, 1),              # WRN-28-10 stage 2: BLOCK_N = 160 of 320 columns, 32-channel dY boxes
    (2, 8, 8, 640, 640, 3, 1, 1),                # WRN-28-10 stage 3
]


@pytest.mark.parametrize("case", S2_CASES)
def test_convtc_stride2_fprop_and_wgrad(case):
    _run_fprop_wgrad(case)


@pytest.mark.parametrize("case", [c + ((c[1], c[2]),) for c in WGRAD_CASES])
def test_convtc_wgrad(case):
    _run_fprop_wgrad(case, fprop=False)


WGRAD_PAIR_CASES = [
    # N, H, W, Cin, Cout, k, stride, pad, (Ho, Wo)
    (8, 14, 14, 256, 256, 3, 1, 1, (14, 14)),     # 36 row blocks: 9 row groups -> 5 pairs, the last one half empty
    (8, 14, 14, 256, 1024, 1, 1, 0, (14, 14)),    # one row group: the pair splits the CTA's two accumulators
    (8, 14, 14, 1024, 256, 1, 1, 0, (14, 14)),
    (4, 28, 28, 128, 128, 3, 1, 1, (28, 28)),     # BLOCK_N = 128: each CTA loads one 64-channel dY box
    (4, 28, 28, 256, 256, 3, 2, 1, (14, 14)),     # stride 2
    (2, 7, 7, 512, 2048, 1, 1, 0, (7, 7)),        # eight column tiles
    (3, 9, 11, 192, 128, 3, 1, 1, (9, 11)),       # ragged pixel tiles, 27 row blocks
    (4, 16, 16, 160, 320, 3, 1, 1, (16, 16)),     # BLOCK_N = 160 is five 32-channel boxes: cannot pair, must fall back
]


@pytest.mark.parametrize("case", WGRAD_PAIR_CASES)
def test_convtc_wgrad_cta_pairs(case, monkeypatch):
    """wgrad as clusters of two CTAs (tcgen05.mma.cta_group::2, M = 256; RN_WG_PAIR=1 forces it wherever the dY tile
    splits in two) against the oracle, and against the single-CTA kernel (same products, fp32 sums in another order)."""
    from distributed_tensorflow_resnet_b200._lib import describe_plan
    N, H, W, Cin, Cout, k, s, pad, (Ho, Wo) = case
    g = torch.Generator().manual_seed(sum(case[:8]) + 5)
    x = torch.randn(N, H, W, Cin, generator=g).to(BF)
    w = (torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5).to(BF)
    dy = torch.randn(N, Ho, Wo, Cout, generator=g).to(BF)
    net = ro.Net(dtype=torch.float64, params={"w": w.double().numpy()})
    xv = ro.Var(x.double())
    yv = net.conv2d(xv, "w", k, Cin, Cout, s, pad, (Ho, Wo), None, "y")
    yv.grad = dy.double()
    net.backward()
    want = torch.as_tensor(net.param_grads["w"])
    d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, k, s, pad, RN_BF16)
    X, DY = dev(x), dev(dy)
    got = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("RN_WG_PAIR", mode)
        nb = lib.rn_convtc_scratch_bytes(C.byref(d), 2)
        # scratch and dW sit between 4 KB guard zones that must come back untouched (the epilogue stores through
        # computed addresses, not through a bounds-checked tensor map)
        G = 4096
        sbuf = torch.full((max(nb, 16) + 2 * G,), 0xA5, dtype=torch.uint8, device="cuda")
        scratch = sbuf[G:G + max(nb, 16)]
        n_dw = k * k * Cin * Cout
        dbuf = torch.full((n_dw + 2 * G // 4,), float("nan"), dtype=torch.float32, device="cuda")
        dbuf[:G // 4] = 12345.0; dbuf[G // 4 + n_dw:] = 12345.0
        DW = dbuf[G // 4:G // 4 + n_dw].view(k, k, Cin, Cout)
        plan = C.c_void_p()
        check(lib.rn_convtc_plan_create(C.byref(d), 2, ptr(X), ptr(DY), ptr(DW), ptr(scratch), nb, C.byref(plan)), "plan wgrad")
        desc = describe_plan(plan)
        if mode == "1" and Cout != 320:
            assert "pair=1" in desc, desc
        else:
            assert "pair=0" in desc, desc
        for _ in range(2):
            check(lib.rn_convtc_run(plan, 0.0, stream()), "run wgrad")
        torch.cuda.synchronize()
        lib.rn_convtc_plan_destroy(plan)
        assert torch.isfinite(DW).all(), desc
        assert bool((sbuf[:G] == 0xA5).all()) and bool((sbuf[G + max(nb, 16):] == 0xA5).all()), "scratch guard overwritten: " + desc
        assert bool((dbuf[:G // 4] == 12345.0).all()) and bool((dbuf[G // 4 + n_dw:] == 12345.0).all()), "dW guard overwritten: " + desc
        assert rel_err(DW.double().cpu(), want) < 2e-3, desc
        got[mode] = DW.clone()
    assert rel_err(got["1"].double().cpu(), got["0"].double().cpu()) < 1e-5


@pytest.mark.parametrize("beta", [0.0, 1.0])
@pytest.mark.parametrize("case", S2_CASES + [c + ((c[1], c[2]),) for c in (CASES[3], CASES[5], CASES[7])])
def test_convtc_dgrad_stride2_and_accumulate(case, beta):
    """dgrad at stride 2 (one GEMM per output parity class through strided views of dx) and
    gradient fan-in (dx += ..., done by the TMA unit as a bf16 reduce-add)."""
    N, H, W, Cin, Cout, k, s, pad, (Ho, Wo) = case
    g = torch.Generator().manual_seed(sum(case[:8]) + 1)
    w = (torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5).to(BF)
    dy = torch.randn(N, Ho, Wo, Cout, generator=g).to(BF)
    base = torch.randn(N, H, W, Cin, generator=g).to(BF)
    net = ro.Net(dtype=torch.float64, params={"w": w.double().numpy()})
    xv = ro.Var(torch.zeros(N, H, W, Cin, dtype=torch.float64))
    yv = net.conv2d(xv, "w", k, Cin, Cout, s, pad, (Ho, Wo), None, "y")
    yv.grad = dy.double()
    net.backward()
    want = xv.grad + (base.double() if beta else 0.0)
    d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, k, s, pad, RN_BF16)
    assert lib.rn_convtc_supported(C.byref(d), 1) == 1
    _, wd = _weights(w)
    DY = dev(dy)
    DX = dev(base) if beta else torch.full((N, H, W, Cin), float("nan"), dtype=BF, device="cuda")
    plan = C.c_void_p()
    check(lib.rn_convtc_plan_create(C.byref(d), 1, ptr(DY), ptr(wd), ptr(DX), None, 0, C.byref(plan)), "plan")
    check(lib.rn_convtc_run(plan, beta, stream()), "run")
    torch.cuda.synchronize()
    lib.rn_convtc_plan_destroy(plan)
    assert torch.isfinite(DX.float()).all()
    assert rel_err(DX.float(), want) < 2e-2


def _run_fprop_wgrad(case, fprop=True):
    N, H, W, Cin, Cout, k, s, pad, (Ho, Wo) = case
    g = torch.Generator().manual_seed(sum(case[:8]))
    x = torch.randn(N, H, W, Cin, generator=g).to(BF)
    w = (torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5).to(BF)
    dy = torch.randn(N, Ho, Wo, Cout, generator=g).to(BF)
    net = ro.Net(dtype=torch.float64, params={"w": w.double().numpy()})
    xv = ro.Var(x.double())
    yv = net.conv2d(xv, "w", k, Cin, Cout, s, pad, (Ho, Wo), None, "y")
    yv.grad = dy.double()
    net.backward()
    d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, k, s, pad, RN_BF16)
    X, DY = dev(x), dev(dy)
    if fprop:
        assert lib.rn_convtc_supported(C.byref(d), 0) == 1
        _, wd = _weights(w)
        Y = torch.full((N, Ho, Wo, Cout), float("nan"), dtype=BF, device="cuda")
        plan = C.c_void_p()
        check(lib.rn_convtc_plan_create(C.byref(d), 0, ptr(X), ptr(wd), ptr(Y), None, 0, C.byref(plan)), "plan")
        check(lib.rn_convtc_run(plan, 0.0, stream()), "run")
        torch.cuda.synchronize()
        lib.rn_convtc_plan_destroy(plan)
        assert torch.isfinite(Y.float()).all()
        assert rel_err(Y.float(), yv.val) < 2e-2
    assert lib.rn_convtc_supported(C.byref(d), 2) == 1
    nb = lib.rn_convtc_scratch_bytes(C.byref(d), 2)
    scratch = torch.empty(max(nb, 16), dtype=torch.uint8, device="cuda")
    DW = torch.full((k, k, Cin, Cout), float("nan"), dtype=torch.float32, device="cuda")
    plan = C.c_void_p()
    check(lib.rn_convtc_plan_create(C.byref(d), 2, ptr(X), ptr(DY), ptr(DW), ptr(scratch), nb, C.byref(plan)), "plan wgrad")
    outs = []
    for _ in range(2):
        check(lib.rn_convtc_run(plan, 0.0, stream()), "run wgrad")
        torch.cuda.synchronize()
        outs.append(DW.clone())
    lib.rn_convtc_plan_destroy(plan)
    assert torch.isfinite(DW).all()
    assert torch.equal(outs[0], outs[1])                        # deterministic split-K
    assert rel_err(DW, net.param_grads["w"]) < 1e-3             # fp32 accumulation of exact bf16 products
    if nb > 0:
        plan = C.c_void_p()
        assert lib.rn_convtc_plan_create(C.byref(d), 2, ptr(X), ptr(DY), ptr(DW), ptr(scratch), nb - 1, C.byref(plan)) == -2


def test_convtc_rejects_unsupported():
    d = ConvDesc(2, 8, 8, 3, 8, 8, 16, 3, 1, 1, RN_BF16)       # 3 input channels: 6-byte rows cannot be TMA'd
    assert lib.rn_convtc_supported(C.byref(d), 0) == 0
    d = ConvDesc(2, 8, 8, 64, 8, 8, 64, 3, 1, 1, 0)            # fp32 verification mode stays on CUDA cores
    assert lib.rn_convtc_supported(C.byref(d), 0) == 0


STEM_S2D_CASES = [
    # N, H, W, Cin, Cout, k, pad, (Ho, Wo), source dtype
    (4, 224, 224, 3, 64, 7, 3, (112, 112), torch.float32),    # ImageNet stem (fixed_padding 3 + VALID)
    (3, 64, 64, 3, 64, 7, 3, (32, 32), BF),
    (2, 37, 41, 3, 64, 7, 2, (19, 21), torch.float32),         # odd sizes, TF-SAME style padding (2 before)
    (2, 32, 32, 4, 32, 8, 3, (16, 16), BF),                    # k = 8, four channels: every packed slot used
]


@pytest.mark.parametrize("case", STEM_S2D_CASES)
def test_stem_space_to_depth_fprop_and_wgrad(case):
    """7x7/2 stem as space-to-depth + 4 vertical taps over a 64-channel overlapping-window view
    (rn_stem_s2d_*), forward and weight gradient, against the oracle's direct convolution."""
    N, H, W, Cin, Cout, k, pad, (Ho, Wo), sdt = case
    g = torch.Generator().manual_seed(sum(case[:7]))
    x = torch.randn(N, H, W, Cin, generator=g).to(BF)
    w = (torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5).to(BF)
    dy = torch.randn(N, Ho, Wo, Cout, generator=g).to(BF)
    net = ro.Net(dtype=torch.float64, params={"w": w.double().numpy()})
    xv = ro.Var(x.double())
    yv = net.conv2d(xv, "w", k, Cin, Cout, 2, pad, (Ho, Wo), None, "y")
    yv.grad = dy.double()
    net.backward()
    d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, k, 2, pad, RN_BF16)
    assert lib.rn_stem_s2d_supported(C.byref(d)) == 1
    H2, W2, T = C.c_int32(), C.c_int32(), C.c_int32()
    check(lib.rn_stem_s2d_geometry(C.byref(d), C.byref(H2), C.byref(W2), C.byref(T)), "geometry")
    assert (H2.value, W2.value, T.value) == (Ho + 3, Wo + 3, 4)
    X = dev(x.to(sdt))
    U = torch.full((N, H2.value, W2.value, 16), float("nan"), dtype=BF, device="cuda")
    check(lib.rn_stem_s2d_pack(C.byref(d), ptr(X), 1 if sdt == BF else 0, ptr(U), stream()), "pack")
    torch.cuda.synchronize()
    # the packing itself is exact: U[n,h2,w2,(dy*2+dx)*Cin+c] = x_padded[n, 2h2+dy, 2w2+dx, c]
    xp = torch.zeros(N, 2 * H2.value, 2 * W2.value, Cin, dtype=BF)
    hh, ww = min(H, 2 * H2.value - pad), min(W, 2 * W2.value - pad)
    xp[:, pad:pad + hh, pad:pad + ww] = x[:, :hh, :ww]
    want_u = torch.zeros(N, H2.value, W2.value, 16, dtype=BF)
    for dy_ in range(2):
        for dx_ in range(2):
            q = (dy_ * 2 + dx_) * Cin
            want_u[..., q:q + Cin] = xp[:, dy_::2, dx_::2]
    assert torch.equal(U.cpu(), want_u)
    WB = dev(w)
    WP = torch.full((4, 64, Cout), float("nan"), dtype=BF, device="cuda")
    check(lib.rn_stem_s2d_weights(C.byref(d), ptr(WB), ptr(WP), stream()), "weights")
    Y = torch.full((N, Ho, Wo, Cout), float("nan"), dtype=BF, device="cuda")
    plan = C.c_void_p()
    check(lib.rn_stem_s2d_plan_create(C.byref(d), 0, ptr(U), ptr(WP), ptr(Y), None, 0, C.byref(plan)), "plan fprop")
    for _ in range(2):
        check(lib.rn_convtc_run(plan, 0.0, stream()), "run")
    torch.cuda.synchronize()
    lib.rn_convtc_plan_destroy(plan)
    assert torch.isfinite(Y.float()).all()
    assert rel_err(Y.float(), yv.val) < 2e-2
    # weight gradient
    DY = dev(dy)
    nb = lib.rn_stem_s2d_scratch_bytes(C.byref(d), 2)
    scratch = torch.zeros(max(nb, 16), dtype=torch.uint8, device="cuda")
    DWP = torch.full((4 * 64, Cout), float("nan"), dtype=torch.float32, device="cuda")
    DW = torch.full((k, k, Cin, Cout), float("nan"), dtype=torch.float32, device="cuda")
    check(lib.rn_stem_s2d_plan_create(C.byref(d), 2, ptr(U), ptr(DY), ptr(DWP), ptr(scratch), nb, C.byref(plan)), "plan wgrad")
    for _ in range(2):
        check(lib.rn_convtc_run(plan, 0.0, stream()), "run")
    check(lib.rn_stem_s2d_wgrad_unpack(C.byref(d), ptr(DWP), ptr(DW), stream()), "unpack")
    torch.cuda.synchronize()
    lib.rn_convtc_plan_destroy(plan)
    assert torch.isfinite(DW).all()
    assert rel_err(DW, net.param_grads["w"]) < 1e-3             # fp32 accumulation of exact bf16 products


FUSED_CASES = [
    # N, H, W, Cin, Cout, k, stride, pad, (Ho, Wo), residual?, stats?
    (4, 16, 16, 64, 64, 3, 1, 1, (16, 16), False, True),       # conv-fed BN: statistics only
    (8, 14, 14, 256, 256, 3, 1, 1, (14, 14), False, True),
    (4, 28, 28, 64, 256, 1, 1, 0, (28, 28), True, True),       # bottleneck conv3 + shortcut + next block's BN statistics
    (6, 7, 7, 512, 2048, 1, 1, 0, (7, 7), True, True),         # stage-4 widths: 8 column tiles, ragged pixel box (98 of 128 rows)
    (3, 9, 11, 64, 128, 1, 1, 0, (9, 11), True, False),        # add only, ragged everything
    (2, 56, 56, 64, 256, 1, 1, 0, (56, 56), True, True),       # C3 stage-1 shape
    (4, 16, 16, 64, 128, 3, 2, 1, (8, 8), False, True),        # stride 2
    (2, 14, 14, 256, 256, 3, 1, 1, (14, 14), False, True),     # box 14x4x2: tile rows past the image bottom are non-zero for k = 3
    (2, 28, 28, 256, 256, 3, 2, 1, (14, 14), False, True),
    (3, 9, 11, 64, 64, 3, 1, 1, (9, 11), True, True),
    (8, 32, 32, 16, 16, 3, 1, 1, (32, 32), False, True),       # CIFAR widths: 16-channel store groups (statistics only)
    (8, 32, 32, 16, 16, 3, 1, 1, (32, 32), True, True),        # basic block's conv2 + shortcut + next BN's statistics
    (8, 16, 16, 32, 32, 3, 1, 1, (16, 16), True, True),        # 32-channel store groups
    (4, 32, 32, 160, 160, 3, 1, 1, (32, 32), True, True),      # WRN-28-10 width: N = 160 in five 32-channel groups
    (8, 32, 32, 16, 32, 3, 2, 1, (16, 16), False, True),       # stage transition
    (3, 9, 11, 32, 48, 3, 1, 1, (9, 11), False, True),         # N = 48: three 16-channel groups, ragged tiles
]


@pytest.mark.parametrize("case", FUSED_CASES)
def test_convtc_fused_epilogue(case):
    """fprop with the residual add and / or the batch-norm statistics fused into the epilogue
    (rn_convtc_plan_create_fused + rn_bn_finalize) against conv -> add -> statistics of the oracle.
    The fused output must equal bf16(bf16(conv) + residual) — the value the unfused pipeline stores —
    and the statistics must be those of the stored tensor."""
    N, H, W, Cin, Cout, k, s, pad, (Ho, Wo), use_res, use_stats = case
    g = torch.Generator().manual_seed(sum(case[:8]) + 7)
    x = torch.randn(N, H, W, Cin, generator=g).to(BF)
    w = (torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5).to(BF)
    res = torch.randn(N, Ho, Wo, Cout, generator=g).to(BF)
    net = ro.Net(dtype=torch.float64, params={"w": w.double().numpy()})
    yv = net.conv2d(ro.Var(x.double()), "w", k, Cin, Cout, s, pad, (Ho, Wo), None, "y")
    d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, k, s, pad, RN_BF16)
    assert lib.rn_convtc_fused_supported(C.byref(d)) == 1
    X, WB, R = dev(x), dev(w), dev(res)
    # unfused reference through the library itself: conv -> (add) -> statistics kernel
    Y0 = torch.empty(N, Ho, Wo, Cout, dtype=BF, device="cuda")
    plan = C.c_void_p()
    check(lib.rn_convtc_plan_create(C.byref(d), 0, ptr(X), ptr(WB), ptr(Y0), None, 0, C.byref(plan)), "plan")
    check(lib.rn_convtc_run(plan, 0.0, stream()), "run")
    torch.cuda.synchronize()
    lib.rn_convtc_plan_destroy(plan)
    want = (Y0.float() + R.float()).to(BF) if use_res else Y0
    rows_cap = lib.rn_convtc_stat_rows(C.byref(d))
    PART = torch.full((rows_cap, Cout, 2), float("nan"), dtype=torch.float32, device="cuda")
    Z = torch.full((N, Ho, Wo, Cout), float("nan"), dtype=BF, device="cuda")
    nrows = C.c_int32(0)
    check(lib.rn_convtc_plan_create_fused(C.byref(d), ptr(X), ptr(WB), ptr(Z), ptr(R) if use_res else None,
                                          ptr(PART) if use_stats else None, C.byref(nrows), C.byref(plan)), "plan fused")
    for _ in range(2):
        check(lib.rn_convtc_run(plan, 0.0, stream()), "run fused")
    torch.cuda.synchronize()
    assert torch.isfinite(Z.float()).all()
    assert torch.equal(Z, want), "fused output differs from bf16(bf16(conv) + residual): max |d| = %g" % (
        (Z.float() - want.float()).abs().max().item())
    assert rel_err(Z.float(), yv.val + (res.double() if use_res else 0.0)) < 2e-2
    assert lib.rn_convtc_run(plan, 1.0, stream()) != 0          # a fused plan cannot accumulate
    if use_stats:
        # folded finalisation: the last CTA of the SAME plan writes mean / rstd / moving averages (ticket resets itself)
        M = N * Ho * Wo
        f_mean = torch.full((Cout,), float("nan"), device="cuda"); f_rstd = torch.full((Cout,), float("nan"), device="cuda")
        f_mm = torch.zeros(Cout, device="cuda"); f_mv = torch.ones(Cout, device="cuda")
        ticket = torch.zeros(1, dtype=torch.int32, device="cuda")
        check(lib.rn_convtc_plan_fold_bn_finalize(plan, M, 1e-5, 0.9, 1, ptr(f_mean), ptr(f_rstd), ptr(f_mm), ptr(f_mv),
                                                  ptr(ticket)), "fold")
        check(lib.rn_convtc_run(plan, 0.0, stream()), "run folded")
        torch.cuda.synchronize()
        assert int(ticket.item()) == 0 and torch.equal(Z, want)
        fold1 = [t.clone() for t in (f_mean, f_rstd, f_mm, f_mv)]
        f_mean.fill_(float("nan")); f_rstd.fill_(float("nan"))
        check(lib.rn_convtc_run(plan, 0.0, stream()), "run folded again")
        torch.cuda.synchronize()
        assert int(ticket.item()) == 0 and torch.equal(f_mean, fold1[0]) and torch.equal(f_rstd, fold1[1])
    else:
        ticket = torch.zeros(1, dtype=torch.int32, device="cuda")
        assert lib.rn_convtc_plan_fold_bn_finalize(plan, 1, 1e-5, 0.9, 1, ptr(ticket), ptr(ticket), None, None, ptr(ticket)) != 0
    lib.rn_convtc_plan_destroy(plan)
    if use_stats:
        assert 0 < nrows.value <= rows_cap and torch.isfinite(PART[:nrows.value]).all()
        mean = torch.empty(Cout, device="cuda"); rstd = torch.empty(Cout, device="cuda")
        mm = torch.zeros(Cout, device="cuda"); mv = torch.ones(Cout, device="cuda")
        check(lib.rn_bn_finalize(ptr(PART), nrows.value, M, Cout, 1e-5, 0.9, 1, ptr(mean), ptr(rstd), ptr(mm), ptr(mv),
                                 stream()), "finalize")
        torch.cuda.synchronize()
        z = Z.double().reshape(M, Cout).cpu()
        m_ref, v_ref = z.mean(0), z.var(0, unbiased=False)
        assert (mean.cpu().double() - m_ref).abs().max() < 1e-5 * max(1.0, float(m_ref.abs().max()))
        assert rel_err(rstd.cpu().double(), 1.0 / torch.sqrt(v_ref + 1e-5)) < 1e-5
        assert rel_err(mm.cpu().double(), 0.1 * m_ref) < 1e-4
        assert rel_err(mv.cpu().double(), 0.9 + 0.1 * v_ref * M / (M - 1)) < 1e-5
        # the folded finalisation sums the same rows in fp64 (another fixed order): equal to the standalone kernel
        for got, ref in zip(fold1, (mean, rstd, mm, mv)):
            assert rel_err(got.cpu().double(), ref.cpu().double()) < 1e-6
        # consumer-side finalisation: the apply pass sums the rows itself (rn_bn_apply_act_fwd_partials) — same
        # statistics, and its output equals the plain apply pass run with the statistics it wrote, bit for bit
        gam = (1.0 + 0.1 * torch.randn(Cout, generator=g)).float().cuda()
        bet = (0.1 * torch.randn(Cout, generator=g)).float().cuda()
        a_mean = torch.full((Cout,), float("nan"), device="cuda"); a_rstd = torch.full((Cout,), float("nan"), device="cuda")
        a_mm = torch.zeros(Cout, device="cuda"); a_mv = torch.ones(Cout, device="cuda")
        Y1 = torch.full_like(Z, float("nan")); Y0 = torch.full_like(Z, float("nan"))
        check(lib.rn_bn_apply_act_fwd_partials(ptr(Z), ptr(Y1), M, Cout, RN_BF16, ptr(gam), ptr(bet), ptr(PART), nrows.value,
                                               1e-5, 0.9, 1, ptr(a_mean), ptr(a_rstd), ptr(a_mm), ptr(a_mv), 0.0, stream()),
              "apply partials")
        check(lib.rn_bn_apply_act_fwd(ptr(Z), ptr(Y0), M, Cout, RN_BF16, ptr(gam), ptr(bet), ptr(a_mean), ptr(a_rstd), 0,
                                      1e-5, 0.0, stream()), "apply")
        torch.cuda.synchronize()
        for got, ref in zip((a_mean, a_rstd, a_mm, a_mv), (mean, rstd, mm, mv)):
            assert rel_err(got.cpu().double(), ref.cpu().double()) < 1e-6
        assert torch.equal(Y1, Y0)


FUSED_DGRAD_CASES = [
    # N, H, W, Cin, Cout, k, pad  (stride 1; dx has Cin channels = the BN's channels)
    (4, 16, 16, 64, 64, 3, 1),
    (8, 14, 14, 256, 256, 3, 1),
    (2, 14, 14, 256, 64, 3, 1),        # pixel box 14x4x2: tile rows below the image are non-zero for k = 3
    (4, 28, 28, 512, 128, 1, 0),       # dgrad of a bottleneck conv1: wide dx (two column tiles), short reduction
    (3, 9, 11, 64, 128, 1, 0),         # ragged
    (6, 7, 7, 1024, 256, 1, 0),        # four column tiles
    (8, 32, 32, 16, 16, 3, 1),         # CIFAR widths: 16-channel store groups
    (8, 16, 16, 32, 32, 3, 1),         # 32-channel store groups
    (3, 9, 11, 48, 32, 3, 1),          # dx has 48 channels: three 16-channel groups, ragged tiles
]


@pytest.mark.parametrize("leak", [0.0, 0.1])
@pytest.mark.parametrize("case", FUSED_DGRAD_CASES)
def test_convtc_fused_dgrad_bn_backward_sums(case, leak):
    """dgrad with the following batch-norm's backward reductions fused into the epilogue
    (rn_convtc_plan_create_fused_dgrad + rn_bn_bwd_finalize): dx must be bit-identical to the plain dgrad,
    dgamma / dbeta must match rn_bn_act_bwd_reduce run on that dx, and the oracle's BN backward."""
    N, H, W, Cin, Cout, k, pad = case
    g = torch.Generator().manual_seed(sum(case) + 11)
    w = (torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5).to(BF)
    dy = torch.randn(N, H, W, Cout, generator=g).to(BF)
    xbn = torch.randn(N, H, W, Cin, generator=g).to(BF)            # the BN's input (its output fed this conv)
    gamma = (1.0 + 0.1 * torch.randn(Cin, generator=g)).float()
    beta = (0.1 * torch.randn(Cin, generator=g)).float()
    M = N * H * W
    x2 = xbn.double().reshape(M, Cin)
    mean = x2.mean(0).float()
    rstd = (1.0 / torch.sqrt(x2.var(0, unbiased=False) + 1e-5)).float()
    d = ConvDesc(N, H, W, Cin, H, W, Cout, k, 1, pad, RN_BF16)
    assert lib.rn_convtc_fused_dgrad_supported(C.byref(d)) == 1
    DY, WB, XB = dev(dy), dev(w), dev(xbn)
    G_, B_, MU, RS = dev(gamma), dev(beta), dev(mean), dev(rstd)
    DX0 = torch.empty(N, H, W, Cin, dtype=BF, device="cuda")
    plan = C.c_void_p()
    check(lib.rn_convtc_plan_create(C.byref(d), 1, ptr(DY), ptr(WB), ptr(DX0), None, 0, C.byref(plan)), "plan")
    check(lib.rn_convtc_run(plan, 0.0, stream()), "run")
    lib.rn_convtc_plan_destroy(plan)
    nb = lib.rn_bn_scratch_bytes(M, Cin)
    sc = torch.zeros(nb, dtype=torch.uint8, device="cuda")
    DG0, DB0 = torch.empty(Cin, device="cuda"), torch.empty(Cin, device="cuda")
    check(lib.rn_bn_act_bwd_reduce(ptr(DX0), ptr(XB), M, Cin, RN_BF16, ptr(G_), ptr(B_), ptr(MU), ptr(RS), leak,
                                   ptr(DG0), ptr(DB0), ptr(sc), nb, stream()), "reduce")
    rows_cap = lib.rn_convtc_stat_rows(C.byref(d))
    PART = torch.full((rows_cap, Cin, 2), float("nan"), dtype=torch.float32, device="cuda")
    DX = torch.full((N, H, W, Cin), float("nan"), dtype=BF, device="cuda")
    nrows = C.c_int32(0)
    check(lib.rn_convtc_plan_create_fused_dgrad(C.byref(d), ptr(DY), ptr(WB), ptr(DX), ptr(XB), ptr(G_), ptr(B_), ptr(MU),
                                                ptr(RS), leak, ptr(PART), C.byref(nrows), C.byref(plan)), "plan fused")
    for _ in range(2):
        check(lib.rn_convtc_run(plan, 0.0, stream()), "run fused")
    DG, DB = torch.empty(Cin, device="cuda"), torch.empty(Cin, device="cuda")
    check(lib.rn_bn_bwd_finalize(ptr(PART), nrows.value, Cin, ptr(DG), ptr(DB), stream()), "finalize")
    torch.cuda.synchronize()
    assert lib.rn_convtc_run(plan, 1.0, stream()) != 0
    # folded finalisation: dgamma / dbeta written by the dgrad's last CTA
    FDG, FDB = torch.full((Cin,), float("nan"), device="cuda"), torch.full((Cin,), float("nan"), device="cuda")
    ticket = torch.zeros(1, dtype=torch.int32, device="cuda")
    check(lib.rn_convtc_plan_fold_bn_bwd_finalize(plan, ptr(FDG), ptr(FDB), ptr(ticket)), "fold")
    for _ in range(2):
        check(lib.rn_convtc_run(plan, 0.0, stream()), "run folded")
    torch.cuda.synchronize()
    assert int(ticket.item()) == 0
    assert rel_err(FDG.cpu().double(), DG.cpu().double()) < 1e-6 and rel_err(FDB.cpu().double(), DB.cpu().double()) < 1e-6
    lib.rn_convtc_plan_destroy(plan)
    assert torch.equal(DX, DX0)
    # consumer-side finalisation: rn_bn_act_bwd_apply_partials sums the rows itself, writes dgamma / dbeta, and its dx
    # equals the plain pass run with those sums (with and without an addend)
    for use_add in (False, True):
        ADDEND = torch.randn(N, H, W, Cin, generator=g).to(BF).cuda() if use_add else None
        ADG, ADB = torch.full((Cin,), float("nan"), device="cuda"), torch.full((Cin,), float("nan"), device="cuda")
        O1 = torch.full((N, H, W, Cin), float("nan"), dtype=BF, device="cuda"); O0 = torch.full_like(O1, float("nan"))
        check(lib.rn_bn_act_bwd_apply_partials(ptr(DX), ptr(XB), ptr(O1), ptr(ADDEND) if use_add else None, M, Cin, RN_BF16,
                                               ptr(G_), ptr(B_), ptr(MU), ptr(RS), ptr(PART), nrows.value, ptr(ADG), ptr(ADB),
                                               leak, stream()), "bwd apply partials")
        check(lib.rn_bn_act_bwd_apply(ptr(DX), ptr(XB), ptr(O0), ptr(ADDEND) if use_add else None, M, Cin, RN_BF16, ptr(G_),
                                      ptr(B_), ptr(MU), ptr(RS), ptr(ADG), ptr(ADB), leak, stream()), "bwd apply")
        torch.cuda.synchronize()
        assert rel_err(ADG.cpu().double(), DG.cpu().double()) < 1e-6 and rel_err(ADB.cpu().double(), DB.cpu().double()) < 1e-6
        assert torch.equal(O1, O0)
    # against the separate reduction kernel (same inputs, different summation order)
    gd = DX0.double().reshape(M, Cin).cpu()
    pre = torch.addcmul((beta - mean * (gamma * rstd)).float(), xbn.float().reshape(M, Cin), (gamma * rstd).float()).double()
    gm = torch.where(pre < 0, leak * gd, gd)
    xhat = (x2 - mean.double()) * rstd.double()
    sb_ref, sg_ref = gm.sum(0), (gm * xhat).sum(0)
    scale_b = max(float(gm.abs().sum(0).max()), 1e-30)
    scale_g = max(float((gm * xhat).abs().sum(0).max()), 1e-30)
    assert (DB.cpu().double() - sb_ref).abs().max().item() / scale_b < 1e-5
    assert (DG.cpu().double() - sg_ref).abs().max().item() / scale_g < 1e-5
    assert (DB - DB0).abs().max().item() / scale_b < 1e-5 and (DG - DG0).abs().max().item() / scale_g < 1e-5


BNIN_CASES = [
    # N, H, W, Cin, Cout, stride, fused residual?, fused stats?
    (4, 16, 16, 64, 64, 1, False, False),
    (4, 28, 28, 256, 64, 1, False, True),        # conv1 of a bottleneck block (statistics of the next BN fused)
    (4, 28, 28, 64, 256, 1, True, True),         # conv3: shortcut add + statistics
    (2, 56, 56, 256, 512, 2, False, False),      # projection shortcut, stride 2 (parity view)
    (3, 9, 11, 128, 192, 1, False, False),       # ragged pixel tiles, N = 192
    (6, 7, 7, 2048, 512, 1, False, True),        # deepest reduction: 32 k-chunks, 16 KB constants table
    (5, 14, 14, 1024, 256, 1, False, True),
]


@pytest.mark.parametrize("leak", [0.0, 0.1])
@pytest.mark.parametrize("case", BNIN_CASES)
def test_convtc_bn_operand_transform(case, leak, monkeypatch):
    """1x1 convolutions that read the INPUT of a training-mode BN+ReLU and normalise it in shared memory
    (rn_convtc_plan_create_bnin): fprop (+ fused epilogues) and wgrad must be BIT-IDENTICAL to
    rn_bn_apply_act_fwd followed by the ordinary plans — same operand values, same GEMM — and match the oracle."""
    monkeypatch.setenv("RN_TC_PAIR", "0")        # same tiling (hence the same per-CTA statistics rows) in both pipelines
    N, H, W, Cin, Cout, s, use_res, use_stats = case
    Ho, Wo = (H - 1) // s + 1, (W - 1) // s + 1
    g = torch.Generator().manual_seed(sum(case[:6]) + 5)
    x = (torch.randn(N, H, W, Cin, generator=g) * 1.5 + 0.3).to(BF)
    w = (torch.randn(1, 1, Cin, Cout, generator=g) / Cin ** 0.5).to(BF)
    dy = torch.randn(N, Ho, Wo, Cout, generator=g).to(BF)
    res = torch.randn(N, Ho, Wo, Cout, generator=g).to(BF)
    gamma = (1.0 + 0.1 * torch.randn(Cin, generator=g)).float()
    beta = (0.1 * torch.randn(Cin, generator=g)).float()
    M = N * H * W
    x2 = x.double().reshape(M, Cin)
    mean = x2.mean(0).float()
    rstd = (1.0 / torch.sqrt(x2.var(0, unbiased=False) + 1e-5)).float()
    X, WB, DY, R = dev(x), dev(w), dev(dy), dev(res)
    G_, B_, MU, RS = dev(gamma), dev(beta), dev(mean), dev(rstd)
    d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, 1, s, 0, RN_BF16)
    assert lib.rn_convtc_bnin_supported(C.byref(d), 0) == 1 and lib.rn_convtc_bnin_supported(C.byref(d), 2) == 1
    # reference pipeline: bn_apply -> materialised activations -> ordinary plans
    Yact = torch.empty(N, H, W, Cin, dtype=BF, device="cuda")
    check(lib.rn_bn_apply_act_fwd(ptr(X), ptr(Yact), M, Cin, RN_BF16, ptr(G_), ptr(B_), ptr(MU), ptr(RS), 0, 1e-5, leak, stream()), "apply")
    Z0 = torch.full((N, Ho, Wo, Cout), float("nan"), dtype=BF, device="cuda")
    P0 = torch.zeros(lib.rn_convtc_stat_rows(C.byref(d)), Cout, 2, device="cuda")
    plan = C.c_void_p()
    nr0 = C.c_int32(0)
    if use_res or use_stats:
        check(lib.rn_convtc_plan_create_fused(C.byref(d), ptr(Yact), ptr(WB), ptr(Z0), ptr(R) if use_res else None,
                                              ptr(P0) if use_stats else None, C.byref(nr0), C.byref(plan)), "plan")
    else:
        check(lib.rn_convtc_plan_create(C.byref(d), 0, ptr(Yact), ptr(WB), ptr(Z0), None, 0, C.byref(plan)), "plan")
    check(lib.rn_convtc_run(plan, 0.0, stream()), "run")
    lib.rn_convtc_plan_destroy(plan)
    nb = lib.rn_convtc_scratch_bytes(C.byref(d), 2)
    sc = torch.zeros(max(nb, 16), dtype=torch.uint8, device="cuda")
    DW0 = torch.full((1, 1, Cin, Cout), float("nan"), device="cuda")
    check(lib.rn_convtc_plan_create(C.byref(d), 2, ptr(Yact), ptr(DY), ptr(DW0), ptr(sc), nb, C.byref(plan)), "wplan")
    check(lib.rn_convtc_run(plan, 0.0, stream()), "wrun")
    lib.rn_convtc_plan_destroy(plan)
    # operand-transform plans on the BN input itself
    Z = torch.full((N, Ho, Wo, Cout), float("nan"), dtype=BF, device="cuda")
    P1 = torch.zeros_like(P0)
    nr1 = C.c_int32(0)
    check(lib.rn_convtc_plan_create_bnin(C.byref(d), 0, ptr(X), ptr(G_), ptr(B_), ptr(MU), ptr(RS), leak, ptr(WB), ptr(Z),
                                         ptr(R) if use_res else None, ptr(P1) if use_stats else None, C.byref(nr1), None, 0,
                                         C.byref(plan)), "bnin fprop")
    from distributed_tensorflow_resnet_b200._lib import describe_plan
    assert "bn_operand=1" in describe_plan(plan) and "pair=0" in describe_plan(plan)
    for _ in range(2):
        check(lib.rn_convtc_run(plan, 0.0, stream()), "run bnin")
    lib.rn_convtc_plan_destroy(plan)
    DW = torch.full((1, 1, Cin, Cout), float("nan"), device="cuda")
    sc2 = torch.zeros_like(sc)
    check(lib.rn_convtc_plan_create_bnin(C.byref(d), 2, ptr(X), ptr(G_), ptr(B_), ptr(MU), ptr(RS), leak, ptr(DY), ptr(DW), None, None,
                                         None, ptr(sc2), nb, C.byref(plan)), "bnin wgrad")
    assert "bn_operand=1" in describe_plan(plan)
    for _ in range(2):
        check(lib.rn_convtc_run(plan, 0.0, stream()), "wrun bnin")
    lib.rn_convtc_plan_destroy(plan)
    torch.cuda.synchronize()
    assert torch.equal(X.cpu(), x)                                   # the BN input in HBM is untouched
    assert torch.isfinite(Z.float()).all() and torch.equal(Z, Z0)
    assert torch.isfinite(DW).all() and torch.equal(DW, DW0)
    if use_stats:
        assert nr1.value == nr0.value and torch.equal(P1[:nr1.value], P0[:nr0.value])
    # and against the oracle (fp64 BN + activation rounded to bf16 where the kernel rounds, conv, add)
    pre = (x2 - mean.double()) * rstd.double() * gamma.double() + beta.double()
    act = torch.where(pre < 0, leak * pre, pre).to(BF).double().reshape(N, H, W, Cin)
    net = ro.Net(dtype=torch.float64, params={"w": w.double().numpy()})
    av = ro.Var(act)
    yv = net.conv2d(av, "w", 1, Cin, Cout, s, 0, (Ho, Wo), None, "y")
    yv.grad = dy.double()
    net.backward()
    assert rel_err(Z.float(), yv.val + (res.double() if use_res else 0.0)) < 2e-2
    assert rel_err(DW, net.param_grads["w"]) < 2e-3
    # an unsupported request is refused, not mis-computed
    d3 = ConvDesc(N, H, W, Cin, H, W, Cout, 3, 1, 1, RN_BF16)
    assert lib.rn_convtc_bnin_supported(C.byref(d3), 0) == 0
