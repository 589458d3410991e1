"""ctypes binding of librn_b200.so (the C ABI declared in include/rn_b200.h).

There is NO fallback: if the shared library cannot be built or loaded, importing this module
raises, and every compute call raises on a non-zero status with the library's own message.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "librn_b200.so")

RN_F32, RN_BF16 = 0, 1


class RnError(RuntimeError):
    pass


class ConvDesc(C.Structure):
    """rn_conv_desc (include/rn_b200.h)."""
    _fields_ = [(n, C.c_int32) for n in
                ("N", "H", "W", "Cin", "Ho", "Wo", "Cout", "k", "stride", "pad", "dtype")]


ABI_VERSION = 5           # must equal RN_ABI_VERSION of include/rn_b200.h the .so was compiled from


def _load():
    """Build (a digest compare of csrc/ + the header when nothing changed) and load the library.
    One process builds at a time: ranks started together by torchrun serialise on a lock file, the
    first one compiles, the others find the stamp up to date."""
    import fcntl
    import importlib.util
    spec = importlib.util.spec_from_file_location("_rn_build", os.path.join(HERE, "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    try:
        lock = open(os.path.join(HERE, ".build.lock"), "w")
    except OSError:
        lock = None                       # read-only tree: use what is there
    try:
        if lock is not None:
            fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            mod.build(verbose=False)      # raises if sources changed and nvcc is missing / fails
        except Exception as e:
            if not os.path.exists(SO_PATH):
                raise RnError("librn_b200.so is missing and could not be built (%s); there is no CPU fallback" % e)
            raise
    finally:
        if lock is not None:
            fcntl.flock(lock, fcntl.LOCK_UN)
            lock.close()
    if not os.path.exists(SO_PATH):
        raise RnError("librn_b200.so is missing and could not be built; there is no CPU fallback")
    so = C.CDLL(SO_PATH)
    so.rn_abi_version.restype = C.c_int
    got = so.rn_abi_version()
    if got != ABI_VERSION:
        raise RnError("librn_b200.so reports ABI version %d, these bindings expect %d: rebuild with "
                      "`python distributed-tensorflow-resnet_b200/build.py --force`" % (got, ABI_VERSION))
    return so


_so = _load()

vp, i32, i64, f32, sz, u32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t, C.c_uint32
pdesc = C.POINTER(ConvDesc)

# name -> (restype, argtypes); every symbol declared in include/rn_b200.h must appear here
SIGNATURES = {
    "rn_last_error": (C.c_char_p, []),
    "rn_abi_version": (C.c_int, []),
    "rn_launch_count": (C.c_uint64, []),
    "rn_conv2d_fprop": (C.c_int, [pdesc, vp, vp, vp, vp]),
    "rn_conv2d_dgrad": (C.c_int, [pdesc, vp, vp, vp, f32, vp]),
    "rn_conv2d_wgrad_scratch_bytes": (sz, [pdesc]),
    "rn_conv2d_wgrad": (C.c_int, [pdesc, vp, vp, vp, vp, sz, vp]),
    "rn_convtc_supported": (C.c_int, [pdesc, C.c_int]),
    "rn_convtc_scratch_bytes": (sz, [pdesc, C.c_int]),
    "rn_convtc_plan_create": (C.c_int, [pdesc, C.c_int, vp, vp, vp, vp, sz, C.POINTER(vp)]),
    "rn_convtc_fused_supported": (C.c_int, [pdesc]),
    "rn_convtc_stat_rows": (C.c_int, [pdesc]),
    "rn_convtc_plan_create_fused": (C.c_int, [pdesc, vp, vp, vp, vp, vp, C.POINTER(i32), C.POINTER(vp)]),
    "rn_bn_finalize": (C.c_int, [vp, i32, C.c_int64, i32, f32, f32, i32, vp, vp, vp, vp, vp]),
    "rn_convtc_fused_dgrad_supported": (C.c_int, [pdesc]),
    "rn_convtc_plan_create_fused_dgrad": (C.c_int, [pdesc, vp, vp, vp, vp, vp, vp, vp, vp, f32, vp, C.POINTER(i32),
                                                    C.POINTER(vp)]),
    "rn_bn_bwd_finalize": (C.c_int, [vp, i32, i32, vp, vp, vp]),
    "rn_bn_apply_act_fwd_partials": (C.c_int, [vp, vp, C.c_int64, i32, i32, vp, vp, vp, i32, f32, f32, i32, vp, vp, vp, vp, f32, vp]),
    "rn_bn_act_bwd_apply_partials": (C.c_int, [vp, vp, vp, vp, C.c_int64, i32, i32, vp, vp, vp, vp, vp, i32, vp, vp, f32, vp]),
    "rn_convtc_plan_fold_bn_finalize": (C.c_int, [vp, C.c_int64, f32, f32, i32, vp, vp, vp, vp, vp]),
    "rn_convtc_plan_fold_bn_bwd_finalize": (C.c_int, [vp, vp, vp, vp]),
    "rn_convtc_bnin_supported": (C.c_int, [pdesc, C.c_int]),
    "rn_convtc_plan_create_bnin": (C.c_int, [pdesc, C.c_int, vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, C.POINTER(i32), vp, sz,
                                             C.POINTER(vp)]),
    "rn_convtc_run": (C.c_int, [vp, f32, vp]),
    "rn_convtc_plan_destroy": (C.c_int, [vp]),
    "rn_convtc_plan_describe": (C.c_int, [vp, C.c_char_p, sz]),
    "rn_weights_to_bf16": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, vp, vp, vp]),
    "rn_relayout_job_bytes": (sz, []),
    "rn_weights_to_bf16_batched": (C.c_int, [vp, i32, vp]),
    "rn_im2col": (C.c_int, [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "rn_stem_weights": (C.c_int, [vp, i32, i32, i32, vp, vp]),
    "rn_stem_s2d_supported": (C.c_int, [pdesc]),
    "rn_stem_s2d_geometry": (C.c_int, [pdesc, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "rn_stem_s2d_pack": (C.c_int, [pdesc, vp, i32, vp, vp]),
    "rn_stem_s2d_weights": (C.c_int, [pdesc, vp, vp, vp]),
    "rn_stem_s2d_wgrad_unpack": (C.c_int, [pdesc, vp, vp, vp]),
    "rn_stem_s2d_scratch_bytes": (sz, [pdesc, C.c_int]),
    "rn_stem_s2d_plan_create": (C.c_int, [pdesc, C.c_int, vp, vp, vp, vp, sz, C.POINTER(vp)]),
    "rn_bn_scratch_bytes": (sz, [i64, i32]),
    "rn_bn_stats_fwd": (C.c_int, [vp, vp, vp, i64, i32, i32, f32, f32, i32, vp, vp, vp, vp, vp, sz, vp]),
    "rn_bn_stats_fwd_noalloc": (C.c_int, [vp, vp, vp, i64, i32, i32, f32, f32, i32, vp, vp, vp, vp, vp]),
    "rn_bn_apply_act_fwd": (C.c_int, [vp, vp, i64, i32, i32, vp, vp, vp, vp, i32, f32, f32, vp]),
    "rn_bn_act_bwd_reduce": (C.c_int, [vp, vp, i64, i32, i32, vp, vp, vp, vp, f32, vp, vp, vp, sz, vp]),
    "rn_bn_act_bwd_apply": (C.c_int, [vp, vp, vp, vp, i64, i32, i32, vp, vp, vp, vp, vp, vp, f32, vp]),
    "rn_maxpool3x3s2_fwd": (C.c_int, [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "rn_maxpool3x3s2_bwd": (C.c_int, [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, f32, vp]),
    "rn_avgpool_fwd": (C.c_int, [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "rn_avgpool_bwd": (C.c_int, [vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, f32, vp]),
    "rn_dense_scratch_bytes": (sz, [i32, i32, i32]),
    "rn_dense_fwd": (C.c_int, [vp, vp, vp, vp, i32, i32, i32, vp, sz, vp]),
    "rn_dense_bwd_scratch_bytes": (sz, [i32, i32, i32]),
    "rn_dense_bwd": (C.c_int, [vp, vp, vp, vp, vp, vp, i32, i32, i32, vp, sz, vp]),
    "rn_softmax_xent": (C.c_int, [vp, vp, i32, i32, vp, vp, vp, vp]),
    "rn_optimizer_scratch_bytes": (sz, [i64]),
    "rn_sgd_momentum_fused": (C.c_int, [vp, vp, vp, vp, i64, vp, vp, f32, f32, f32, i32, vp, vp, sz, vp]),
    "rn_l2_sumsq": (C.c_int, [vp, i64, vp, vp, vp, sz, vp]),
    "rn_cifar_decode": (C.c_int, [vp, i64, i32, i32, i32, vp, i32, i32, i32, u32, u32, u32, vp, i32, vp, i32, vp, vp, vp, vp]),
    "rn_vgg_preprocess": (C.c_int, [vp, vp, vp, i32, i32, i32, i32, i32, i32, u32, u32, u32, vp, vp, vp, i32, vp, vp, vp, i32, vp, vp]),
    "rn_cast": (C.c_int, [vp, vp, i64, i32, i32, vp]),
    "rn_add_inplace": (C.c_int, [vp, vp, i64, i32, vp]),
    "rn_add": (C.c_int, [vp, vp, vp, i64, i32, vp]),
    "rn_pad_channels": (C.c_int, [vp, vp, i64, i32, i32, i32, i32, vp]),
    "rn_crc32c": (u32, [vp, sz, u32]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(_so, _name)   # AttributeError here == a declared symbol is not exported
    _fn.restype = _res
    _fn.argtypes = _args

lib = _so


def last_error():
    return _so.rn_last_error().decode(errors="replace")


def check(rc, what=""):
    if rc != 0:
        raise RnError("%s failed (status %d): %s" % (what or "librn_b200 call", rc, last_error()))


def describe_plan(plan):
    buf = C.create_string_buffer(256)
    check(_so.rn_convtc_plan_describe(plan, buf, 256), "rn_convtc_plan_describe")
    return buf.value.decode()


def launch_count():
    return int(_so.rn_launch_count())


def ptr(t):
    """Device (or host) address of a torch tensor, or None."""
    return None if t is None else t.data_ptr()
