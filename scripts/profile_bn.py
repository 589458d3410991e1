"""Time the four BN kernels on one tensor shape (L2 flushed between runs).
usage: python scripts/profile_bn.py M C [bf16|f32]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_resnet_b200._lib import lib, check, ptr, RN_BF16, RN_F32
M, C = int(sys.argv[1]), int(sys.argv[2])
bf = (sys.argv[3] if len(sys.argv) > 3 else "bf16") == "bf16"
dt, rdt, s = (torch.bfloat16, RN_BF16, 2) if bf else (torch.float32, RN_F32, 4)
x = torch.randn(M, C, device="cuda").to(dt); dy = torch.randn(M, C, device="cuda").to(dt)
add = torch.randn(M, C, device="cuda").to(dt); z = torch.empty_like(x); y = torch.empty_like(x); dx = torch.empty_like(x)
g = torch.ones(C, device="cuda"); b = torch.zeros(C, device="cuda"); mean = torch.zeros(C, device="cuda"); rstd = torch.ones(C, device="cuda")
dg = torch.zeros(C, device="cuda"); db = torch.zeros(C, device="cuda")
nb = lib.rn_bn_scratch_bytes(M, C); sc = torch.zeros(nb, dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
E = M * C * s
def t(fn, nbytes, name):
    best = 1e9
    for _ in range(5):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("  %-16s %8.4f ms  %7.1f GB/s" % (name, best, nbytes / best / 1e6))
print("M=%d C=%d %s (%.1f MB)" % (M, C, "bf16" if bf else "f32", E / 1e6))
t(lambda: check(lib.rn_bn_stats_fwd(ptr(x), None, None, M, C, rdt, 1e-5, 0.9, 1, ptr(mean), ptr(rstd), None, None, ptr(sc), nb, st)), E, "stats")
t(lambda: check(lib.rn_bn_stats_fwd(ptr(x), ptr(add), ptr(z), M, C, rdt, 1e-5, 0.9, 1, ptr(mean), ptr(rstd), None, None, ptr(sc), nb, st)), 3 * E, "stats+add")
t(lambda: check(lib.rn_bn_apply_act_fwd(ptr(x), ptr(y), M, C, rdt, ptr(g), ptr(b), ptr(mean), ptr(rstd), 0, 1e-5, 0.0, st)), 2 * E, "apply")
t(lambda: check(lib.rn_bn_act_bwd_reduce(ptr(dy), ptr(x), M, C, rdt, ptr(g), ptr(b), ptr(mean), ptr(rstd), 0.0, ptr(dg), ptr(db), ptr(sc), nb, st)), 2 * E, "bwd_reduce")
t(lambda: check(lib.rn_bn_act_bwd_apply(ptr(dy), ptr(x), ptr(dx), None, M, C, rdt, ptr(g), ptr(b), ptr(mean), ptr(rstd), ptr(dg), ptr(db), 0.0, st)), 3 * E, "bwd_apply")
