import sys, os, ctypes as C
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
import torch
from util import dev, stream, lib, check, ptr, ConvDesc, RN_BF16
BF = torch.bfloat16
def run(case):
    N, H, W, Cin, Cout, k, s, pad, (Ho, Wo) = case
    g = torch.Generator().manual_seed(1)
    x = torch.randn(N, H, W, Cin, generator=g).to(BF)
    w = (torch.randn(k, k, Cin, Cout, generator=g) / (k * k * Cin) ** 0.5).to(BF)
    d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, k, s, pad, RN_BF16)
    X, WB = dev(x), dev(w)
    rows_cap = lib.rn_convtc_stat_rows(C.byref(d))
    PART = torch.full((rows_cap, Cout, 2), float("nan"), dtype=torch.float32, device="cuda")
    Z = torch.full((N, Ho, Wo, Cout), float("nan"), dtype=BF, device="cuda")
    nrows = C.c_int32(0); plan = C.c_void_p()
    check(lib.rn_convtc_plan_create_fused(C.byref(d), ptr(X), ptr(WB), ptr(Z), None, ptr(PART), C.byref(nrows), C.byref(plan)), "plan")
    check(lib.rn_convtc_run(plan, 0.0, stream()), "run"); torch.cuda.synchronize()
    M = N * Ho * Wo
    z = Z.double().reshape(M, Cout)
    S = PART[:nrows.value, :, 0].double().sum(0); Q = PART[:nrows.value, :, 1].double().sum(0)
    es = (S - z.sum(0)).abs().max().item() / z.abs().sum(0).max().item()
    eq = (Q - (z * z).sum(0)).abs().max().item() / (z * z).sum(0).max().item()
    print(case, "rows", nrows.value, "err sum %.2e sumsq %.2e" % (es, eq))
    if es > 1e-4:
        per = (S - z.sum(0)).abs().cpu()
        print("  bad channels:", (per > 1e-3 * z.abs().sum(0).max().item()).nonzero().flatten()[:40].tolist())
        # per-tile check
        for r in range(nrows.value):
            print("  row", r, "sum of partial", PART[r, :4, 0].tolist())
for c in [(2, 28, 28, 256, 256, 3, 2, 1, (14, 14)), (2, 14, 14, 256, 256, 3, 1, 1, (14, 14)), (2, 14, 14, 64, 64, 1, 1, 0, (14, 14)),
          (2, 14, 14, 64, 256, 1, 1, 0, (14, 14)), (4, 14, 14, 64, 256, 1, 1, 0, (14, 14)), (2, 7, 7, 64, 64, 1, 1, 0, (7, 7))]:
    run(c)
