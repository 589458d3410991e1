"""Time the classifier-head kernels (dense fwd / bwd, softmax cross-entropy).
usage: python scripts/profile_head.py [B D K]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_resnet_b200._lib import lib, check, ptr
B, D, K = [int(a) for a in sys.argv[1:4]] if len(sys.argv) > 3 else (128, 2048, 1001)
x = torch.randn(B, D, device="cuda"); w = torch.randn(D, K, device="cuda") * 0.02; b = torch.zeros(K, device="cuda")
y = torch.empty(B, K, device="cuda"); oh = torch.zeros(B, K, device="cuda"); oh[torch.arange(B), torch.arange(B) % K] = 1
p = torch.empty(B, K, device="cuda"); dl = torch.empty(B, K, device="cuda"); sc2 = torch.empty(2, device="cuda")
dw = torch.empty(D, K, device="cuda"); db = torch.empty(K, device="cuda"); dx = torch.empty(B, D, device="cuda")
nf = lib.rn_dense_scratch_bytes(B, D, K); sf = torch.zeros(max(nf, 16), dtype=torch.uint8, device="cuda")
nb = lib.rn_dense_bwd_scratch_bytes(B, D, K); sb = torch.zeros(max(nb, 16), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def t(fn, name):
    best = 1e9
    for _ in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("  %-22s %8.4f ms" % (name, best))
t(lambda: check(lib.rn_dense_fwd(ptr(x), ptr(w), ptr(b), ptr(y), B, D, K, ptr(sf), nf, st)), "dense fwd (split-K)")
t(lambda: check(lib.rn_softmax_xent(ptr(y), ptr(oh), B, K, ptr(p), ptr(dl), ptr(sc2), st)), "softmax_xent")
t(lambda: check(lib.rn_dense_bwd(ptr(x), ptr(w), ptr(dl), None, ptr(db), ptr(dx), B, D, K, ptr(sb), nb, st)), "dense bwd dx+db splitK")
t(lambda: check(lib.rn_dense_bwd(ptr(x), ptr(w), ptr(dl), None, None, ptr(dx), B, D, K, None, 0, st)), "dense bwd dx one pass")
t(lambda: check(lib.rn_dense_bwd(ptr(x), ptr(w), ptr(dl), ptr(dw), None, None, B, D, K, None, 0, st)), "dense bwd dw")
