"""Run one tensor-core conv GEMM a few times (for ncu / timing).
usage: python scripts/profile_conv.py N H W Cin Cout k stride pad kind [reps]"""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_resnet_b200._lib import lib, check, ptr, ConvDesc, RN_BF16
N, H, W, Cin, Cout, k, s, pad, kind = [int(a) for a in sys.argv[1:10]]
reps = int(sys.argv[10]) if len(sys.argv) > 10 else 5
Ho = H if s == 1 else (H - 1) // s + 1
Wo = W if s == 1 else (W - 1) // s + 1
BF = torch.bfloat16
x = torch.randn(N, H, W, Cin, device="cuda").to(BF)
w = (torch.randn(k, k, Cin, Cout, device="cuda") / (k * k * Cin) ** 0.5).to(BF)
dy = torch.randn(N, Ho, Wo, Cout, device="cuda").to(BF)
y = torch.empty(N, Ho, Wo, Cout, device="cuda", dtype=BF)
dx = torch.empty(N, H, W, Cin, device="cuda", dtype=BF)
dw = torch.empty(k, k, Cin, Cout, device="cuda")
d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, k, s, pad, RN_BF16)
nb = lib.rn_convtc_scratch_bytes(C.byref(d), kind)
scratch = torch.empty(max(nb, 16), dtype=torch.uint8, device="cuda")
a, b, out = [(x, w, y), (dy, w, dx), (x, dy, dw)][kind]
plan = C.c_void_p()
check(lib.rn_convtc_plan_create(C.byref(d), kind, ptr(a), ptr(b), ptr(out), ptr(scratch), nb, C.byref(plan)), "plan")
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ts = []
for i in range(reps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    check(lib.rn_convtc_run(plan, 0.0, st), "run")
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
fl = 2.0 * N * Ho * Wo * Cout * k * k * Cin
print("kind %d  %s: best %.4f ms  %.1f TFLOP/s (L2 flushed between runs)" % (kind, sys.argv[1:9], min(ts), fl / min(ts) / 1e9))
