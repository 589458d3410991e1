"""Run one fprop with the fused epilogue (residual add + BN statistics) a few times (for ncu / timing).
usage: python scripts/profile_fused.py N H W Cin Cout k stride pad [res=1] [stats=1] [reps]"""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_resnet_b200._lib import lib, check, ptr, ConvDesc, RN_BF16
N, H, W, Cin, Cout, k, s, pad = [int(a) for a in sys.argv[1:9]]
use_res = int(sys.argv[9]) if len(sys.argv) > 9 else 1
use_stats = int(sys.argv[10]) if len(sys.argv) > 10 else 1
reps = int(sys.argv[11]) if len(sys.argv) > 11 else 5
Ho = H if s == 1 else (H - 1) // s + 1
Wo = W if s == 1 else (W - 1) // s + 1
BF = torch.bfloat16
x = torch.randn(N, H, W, Cin, device="cuda").to(BF)
w = (torch.randn(k, k, Cin, Cout, device="cuda") / (k * k * Cin) ** 0.5).to(BF)
res = torch.randn(N, Ho, Wo, Cout, device="cuda").to(BF)
z = torch.empty(N, Ho, Wo, Cout, device="cuda", dtype=BF)
d = ConvDesc(N, H, W, Cin, Ho, Wo, Cout, k, s, pad, RN_BF16)
part = torch.zeros(lib.rn_convtc_stat_rows(C.byref(d)), Cout, 2, device="cuda")
plan = C.c_void_p(); nrows = C.c_int32(0)
check(lib.rn_convtc_plan_create_fused(C.byref(d), ptr(x), ptr(w), ptr(z), ptr(res) if use_res else None,
                                      ptr(part) if use_stats else None, C.byref(nrows), C.byref(plan)), "plan")
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ts = []
for i in range(reps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); check(lib.rn_convtc_run(plan, 0.0, st), "run"); e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
byts = (x.numel() + z.numel() * (2 if use_res else 1)) * 2
print("fused res=%d stats=%d %s: best %.4f ms  %.0f GB/s algorithmic" % (use_res, use_stats, sys.argv[1:9], min(ts), byts / min(ts) / 1e6))
