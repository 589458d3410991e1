"""rn_bn_stats_fwd_noalloc: the batch-norm statistics pass for embedders that cannot give the
library any scratch memory (for instance a plugin inside a process whose framework has claimed all
device memory, where cudaMalloc fails).  It keeps its partial sums in a fixed device array inside
the library instead.  The test drives it through the C ABI and requires the same results as the
regular entry point (rn_bn_stats_fwd, caller scratch) on two consecutive calls with different data —
the second call meets the first call's leftovers in the shared array.  Collected last (file name)
so a run that stops at the first failure has already exercised everything else."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from util import rel_err, dev, stream, lib, check, ptr, RN_F32, RN_BF16  # noqa: E402

EPS, DECAY = 1.001e-5, 0.997


def _both(x, addend, M, C, rdt):
    tdt = x.dtype
    outs = []
    for which in ("ref", "noalloc"):
        mean, rstd = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
        mm, mv = torch.full((C,), 0.25, device="cuda"), torch.full((C,), 1.5, device="cuda")
        z = torch.empty(M, C, dtype=tdt, device="cuda") if addend is not None else None
        if which == "ref":
            nb = lib.rn_bn_scratch_bytes(M, C)
            scratch = torch.zeros(nb, dtype=torch.uint8, device="cuda")
            check(lib.rn_bn_stats_fwd(ptr(x), ptr(addend), ptr(z), M, C, rdt, EPS, DECAY, 1, ptr(mean), ptr(rstd),
                                      ptr(mm), ptr(mv), ptr(scratch), nb, stream()), "rn_bn_stats_fwd")
        else:
            check(lib.rn_bn_stats_fwd_noalloc(ptr(x), ptr(addend), ptr(z), M, C, rdt, EPS, DECAY, 1, ptr(mean),
                                              ptr(rstd), ptr(mm), ptr(mv), stream()), "rn_bn_stats_fwd_noalloc")
        torch.cuda.synchronize()
        outs.append((mean, rstd, mm, mv, z))
    return outs


@pytest.mark.parametrize("case", [(128 * 28 * 28, 512, "bf16"), (4 * 32 * 32, 16, "f32"), (98, 2048, "f32")])
def test_noalloc_matches_caller_scratch_entry_on_two_calls(case):
    M, C, mode = case
    tdt, rdt = (torch.bfloat16, RN_BF16) if mode == "bf16" else (torch.float32, RN_F32)
    g = torch.Generator().manual_seed(M + C)
    for call in range(2):       # second call: different data, same static array
        x = dev(torch.randn(M, C, generator=g) * (1.0 + call) + 0.5 * call, tdt)
        addend = dev(torch.randn(M, C, generator=g), tdt) if call == 1 else None
        ref, got = _both(x, addend, M, C, rdt)
        for a, b, what in zip(ref[:4], got[:4], ("mean", "rstd", "moving_mean", "moving_variance")):
            assert rel_err(b, a) < 1e-5, (what, call)
        if addend is not None:
            assert torch.equal(ref[4], got[4])          # the fused sum z = x + addend is bit-identical
        # and directly against the fp64 oracle (FusedBatchNorm semantics, resnet_model_official.py:41-50):
        # statistics of the STORED sum, biased variance in rstd, unbiased variance in the moving average
        z64 = x.double() if addend is None else (x.float() + addend.float()).to(tdt).double()
        mu = z64.mean(0)
        var = z64.var(0, unbiased=False)
        assert rel_err(got[0], mu) < 1e-5 and rel_err(got[1], 1.0 / torch.sqrt(var + EPS)) < 1e-5, call
        assert rel_err(got[2], 0.25 - (0.25 - mu) * (1 - DECAY)) < 1e-5, call
        assert rel_err(got[3], 1.5 - (1.5 - var * M / (M - 1)) * (1 - DECAY)) < 1e-5, call


def test_noalloc_rejects_sizes_beyond_its_limit():
    from distributed_tensorflow_resnet_b200._lib import last_error
    x = torch.zeros(8, 8192, device="cuda")
    m = torch.zeros(8192, device="cuda")
    rc = lib.rn_bn_stats_fwd_noalloc(ptr(x), None, None, 8, 8192, RN_F32, EPS, DECAY, 1, ptr(m), ptr(m), None, None, stream())
    assert rc == -1 and "4096" in last_error()
