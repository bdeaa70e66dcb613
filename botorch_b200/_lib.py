"""ctypes binding of libbotorch_b200.so (the C ABI in include/botorch_b200.h).

There is no fallback: if the library is missing or the device is not a B200-class
GPU every call raises.  PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from .build import LIB_PATH

c_double_p = C.POINTER(C.c_double)


class BoGP(C.Structure):
    _fields_ = [
        ("kernel_id", C.c_int32),
        ("n", C.c_int32),
        ("d", C.c_int32),
        ("M", C.c_int32),
        ("y_mean", C.c_double),
        ("y_std", C.c_double),
        ("train_X", C.c_void_p),
        ("center", C.c_void_p),
        ("inv_ls", C.c_void_p),
        ("outputscale", C.c_void_p),
        ("noise", C.c_void_p),
        ("mean", C.c_void_p),
        ("targets", C.c_void_p),
        ("jitter", C.c_void_p),
        ("aug_train", C.c_void_p),
        ("linv_aug", C.c_void_p),
        ("linv_aug_t", C.c_void_p),
        ("chol", C.c_void_p),
        ("xt_aug", C.c_void_p),
        ("noise_vec", C.c_void_p),
    ]


class BoNEI(C.Structure):
    _fields_ = [
        ("n_b", C.c_int32),
        ("S", C.c_int32),
        ("baseline_is_train", C.c_int32),
        ("X_b", C.c_void_p),
        ("Z_b", C.c_void_p),
        ("jitter", C.c_void_p),
        ("aug_b", C.c_void_p),
        ("vb", C.c_void_p),
        ("vb_t", C.c_void_p),
        ("lb", C.c_void_p),
        ("lb_inv", C.c_void_p),
        ("lb_inv_t", C.c_void_p),
        ("zb_pad", C.c_void_p),
        ("zb_t", C.c_void_p),
        ("mean_b", C.c_void_p),
        ("best_f", C.c_void_p),
        ("qmat", C.c_void_p),
        ("qmat_t", C.c_void_p),
        ("fwd_cat", C.c_void_p),
        ("bwd_cat", C.c_void_p),
        ("reduction", C.c_int32),
        ("no_baseline", C.c_int32),
        ("obj_weight", C.c_double),
        ("external_df", C.c_int32),
        ("schur_jitter", C.c_double),
        ("schur_fixed", C.c_int32),
        ("local_derivs", C.c_int32),
        ("setup_phase", C.c_int32),
    ]


class BoPaths(C.Structure):
    _fields_ = [("S", C.c_int32), ("F", C.c_int32)] + [(name, C.c_void_p) for name in "W w eps w_aug w_pad v w_t v_t".split()]


class BoLBFGS(C.Structure):
    _fields_ = [("B", C.c_int32), ("D", C.c_int32), ("m", C.c_int32)] + [
        (name, C.c_void_p)
        for name in ("x g f x_prev g_prev S Y rho hist_count hist_head p alpha x_trial state lower upper n_iter").split()
    ]


class BoError(RuntimeError):
    pass


_lib: Optional[C.CDLL] = None
_ctx = {}


def _declare(lib: C.CDLL) -> None:
    vp, i32, i64, dbl, sz = C.c_void_p, C.c_int, C.c_long, C.c_double, C.c_size_t
    gp, nei = C.POINTER(BoGP), C.POINTER(BoNEI)
    lib.bo_strerror.restype = C.c_char_p
    lib.bo_strerror.argtypes = [i32]
    lib.bo_last_cuda_error.restype = C.c_char_p
    lib.bo_ctx_create.argtypes = [i32, C.POINTER(vp)]
    lib.bo_ctx_destroy.argtypes = [vp]
    lib.bo_ctx_destroy.restype = None
    lib.bo_launch_count.restype = i64
    lib.bo_launch_count_reset.restype = None
    lib.bo_prof_enable.argtypes = [i32]
    lib.bo_prof_enable.restype = None
    lib.bo_prof_read.argtypes = [C.POINTER(dbl), C.POINTER(i64), C.POINTER(dbl)]
    lib.bo_prof_read.restype = i32
    lib.bo_dmma_peak_launch.argtypes = [i32, vp, C.POINTER(dbl), vp]
    lib.bo_dmma_peak_launch.restype = i32
    lib.bo_source_digest.restype = C.c_char_p
    lib.bo_pad8.restype = i64
    lib.bo_pad8.argtypes = [i64]
    lib.bo_kp.restype = i64
    lib.bo_kp.argtypes = [i32]
    lib.bo_gp_setup_workspace_bytes.restype = sz
    lib.bo_gp_setup_workspace_bytes.argtypes = [gp]
    lib.bo_gp_setup.argtypes = [vp, gp, vp, sz, vp, vp]
    lib.bo_nei_setup_workspace_bytes.restype = sz
    lib.bo_nei_setup_workspace_bytes.argtypes = [gp, nei]
    lib.bo_nei_setup.argtypes = [vp, gp, nei, vp, sz, vp, vp]
    lib.bo_nei_dense_rows.restype = i64
    lib.bo_nei_dense_rows.argtypes = [nei]
    lib.bo_ei_setup_workspace_bytes.restype = sz
    lib.bo_ei_setup_workspace_bytes.argtypes = [gp, nei]
    lib.bo_ei_setup.argtypes = [vp, gp, nei, vp, sz, vp, vp]
    lb = C.POINTER(BoLBFGS)
    lib.bo_lbfgs_init.argtypes = [vp, lb, vp, vp, vp]
    lib.bo_lbfgs_propose.argtypes = [vp, lb, vp]
    lib.bo_lbfgs_accept.argtypes = [vp, lb, vp, vp, dbl, dbl, dbl, i32, vp]
    for name in ("bo_lbfgs_init", "bo_lbfgs_propose", "bo_lbfgs_accept"):
        getattr(lib, name).restype = i32
    lib.bo_qlognei_workspace_bytes.restype = sz
    lib.bo_qlognei_workspace_bytes.argtypes = [gp, nei, i32, i32, i32]
    lib.bo_qlognei_fwd.argtypes = [vp, gp, nei, i32, i32, vp, vp, dbl, dbl, vp, vp, vp, sz, i32, vp]
    lib.bo_qlognei_bwd.argtypes = [vp, gp, nei, i32, i32, vp, vp, dbl, dbl, vp, vp, vp, sz, vp]
    lib.bo_posterior_workspace_bytes.restype = sz
    lib.bo_posterior_workspace_bytes.argtypes = [gp, i64]
    lib.bo_posterior_joint.argtypes = [vp, gp, i32, i32, vp, vp, vp, vp, sz, vp]
    lib.bo_posterior_joint_large_workspace_bytes.restype = sz
    lib.bo_posterior_joint_large_workspace_bytes.argtypes = [gp, i64]
    lib.bo_posterior_joint_large.argtypes = [vp, gp, i64, vp, vp, vp, vp, sz, vp]
    lib.bo_posterior_joint_large.restype = i32
    lib.bo_posterior_bwd_workspace_bytes.restype = sz
    lib.bo_posterior_bwd_workspace_bytes.argtypes = [gp, i32, i32]
    lib.bo_posterior_joint_bwd.argtypes = [vp, gp, i32, i32, vp, vp, vp, vp, vp, sz, vp]
    lib.bo_posterior_mean_var.argtypes = [vp, gp, i64, vp, vp, vp, vp, sz, vp]
    lib.bo_ts_argmax.argtypes = [vp, i32, i64, vp, vp, vp, vp, vp, vp]
    lib.bo_argmax.argtypes = [vp, i64, vp, vp, vp, vp]
    lib.bo_dgemm.argtypes = [vp, i32, i32, i32, i32, dbl, vp, i64, i64, vp, i64, i64, dbl, vp, i64, i64, i32, vp]
    lib.bo_kernel_matrix.argtypes = [vp, gp, i64, vp, vp, vp, sz, vp]
    lib.bo_kernel_matrix_static.argtypes = [vp, gp, i64, vp, vp, vp]
    lib.bo_kernel_matrix_static.restype = i32
    lib.bo_potrf_workspace_bytes.restype = sz
    lib.bo_potrf_workspace_bytes.argtypes = [i32, i32]
    lib.bo_potrf.argtypes = [vp, i32, i32, vp, i64, i64, vp, vp, sz, vp, vp]
    pt = C.POINTER(BoPaths)
    lib.bo_paths_setup_workspace_bytes.restype = sz
    lib.bo_paths_setup_workspace_bytes.argtypes = [gp, pt]
    lib.bo_paths_setup.argtypes = [vp, gp, pt, vp, sz, vp]
    lib.bo_paths_eval_workspace_bytes.restype = sz
    lib.bo_paths_eval_workspace_bytes.argtypes = [gp, pt, i64]
    lib.bo_paths_eval.argtypes = [vp, gp, pt, i64, vp, vp, vp, sz, vp]
    lib.bo_paths_grad_workspace_bytes.restype = sz
    lib.bo_paths_grad_workspace_bytes.argtypes = [gp, pt, i64]
    lib.bo_paths_grad.argtypes = [vp, gp, pt, i64, vp, vp, vp, vp, sz, vp]
    lib.bo_qlognei_samples_view.argtypes = [gp, nei, i32, i32, C.POINTER(i64), C.POINTER(i64)]
    lib.bo_mc_reduce_fwd.argtypes = [vp, i32, i32, i32, i64, vp, vp, vp, dbl, dbl, dbl, i32, vp, vp, vp]
    lib.bo_mc_reduce_bwd.argtypes = [vp, i32, i32, i32, i64, vp, vp, vp, dbl, dbl, dbl, i32, vp, vp, vp, vp, vp]
    lib.bo_mll_workspace_bytes.restype = sz
    lib.bo_mll_workspace_bytes.argtypes = [gp]
    lib.bo_mll_value_grad.argtypes = [vp, gp, vp, sz, vp, vp, vp, vp]
    lib.bo_saas_potential.argtypes = [vp, gp, vp, vp, sz, vp, vp, vp, vp, vp, vp, vp]
    for name in (
        "bo_ctx_create bo_gp_setup bo_nei_setup bo_ei_setup bo_qlognei_fwd bo_qlognei_bwd bo_posterior_joint bo_posterior_joint_bwd "
        "bo_posterior_mean_var bo_ts_argmax bo_argmax bo_dgemm bo_kernel_matrix bo_potrf bo_mll_value_grad bo_saas_potential bo_paths_setup bo_paths_eval bo_paths_grad bo_qlognei_samples_view bo_mc_reduce_fwd bo_mc_reduce_bwd"
    ).split():
        getattr(lib, name).restype = i32


def lib() -> C.CDLL:
    """The loaded shared library; raises if it has not been built (run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise BoError(
                f"{LIB_PATH} not found: the CUDA extension is not built. There is no CPU fallback; "
                "run `python -m botorch_b200.build` (needs nvcc)."
            )
        _lib = C.CDLL(LIB_PATH)
        _declare(_lib)
        _check_fresh(_lib)
        if os.environ.get("BOTORCH_B200_NVTX", "0") == "1":
            _nvtx_wrap(_lib)
    return _lib


def _check_fresh(lib_: C.CDLL) -> None:
    """A binary built from other sources than the ones next to it would run kernels that do not match the code (the
    .sha256 stamp beside the .so is not proof: it can travel through a checkout while the binary stays).  The digest is
    compiled into the library; when the sources are present and differ, refuse to run."""
    from .build import CSRC, SOURCES, source_digest

    if not all(os.path.exists(os.path.join(CSRC, f)) for f in SOURCES):
        return
    built = lib_.bo_source_digest().decode()
    if built not in ("unknown", source_digest()):
        raise BoError(
            f"{LIB_PATH} was built from different sources (digest {built[:12]}..., sources {source_digest()[:12]}...): "
            "run `python -m botorch_b200.build`."
        )


def _nvtx_wrap(lib_: C.CDLL) -> None:
    """BOTORCH_B200_NVTX=1: an NVTX range around every C-ABI call that enqueues work, so that nsys / ncu timelines show
    bo_qlognei_fwd, bo_gp_setup, ... as named host ranges."""
    names = ("bo_gp_setup bo_nei_setup bo_ei_setup bo_qlognei_fwd bo_qlognei_bwd bo_posterior_joint bo_posterior_joint_bwd "
             "bo_posterior_joint_large bo_posterior_mean_var bo_ts_argmax bo_argmax bo_dgemm bo_kernel_matrix bo_potrf "
             "bo_mll_value_grad bo_saas_potential bo_paths_setup bo_paths_eval bo_paths_grad bo_mc_reduce_fwd "
             "bo_mc_reduce_bwd bo_lbfgs_init bo_lbfgs_propose bo_lbfgs_accept bo_kernel_matrix_static").split()  # fmt: skip

    def wrap(name: str, fn):
        def call(*args):
            torch.cuda.nvtx.range_push(name)
            try:
                return fn(*args)
            finally:
                torch.cuda.nvtx.range_pop()

        return call

    for name in names:
        setattr(lib_, name, wrap(name, getattr(lib_, name)))


def check(code: int) -> None:
    if code != 0:
        l = lib()
        msg = l.bo_strerror(code).decode()
        if code == 2:
            msg += ": " + l.bo_last_cuda_error().decode()
        raise BoError(f"libbotorch_b200: {msg}")


def ctx(device: torch.device) -> C.c_void_p:
    """One bo_ctx per device (not thread-safe; the path is driven single-threaded from SciPy's callback)."""
    if device.type != "cuda":
        raise BoError(f"botorch_b200 runs on CUDA devices only (sm_100a); got device {device}. No CPU fallback.")
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _ctx:
        h = C.c_void_p()
        check(lib().bo_ctx_create(idx, C.byref(h)))
        _ctx[idx] = h
    return _ctx[idx]


def stream_ptr(device: torch.device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous() and t.dtype in (torch.float64, torch.int32, torch.int64), (
        t.device,
        t.dtype,
        t.is_contiguous(),
    )
    return t.data_ptr()


def pad8(x: int) -> int:
    return (x + 7) // 8 * 8


def kp(d: int) -> int:
    return (d + 2 + 3) // 4 * 4


def workspace(nbytes: int, device: torch.device) -> torch.Tensor:
    """Caller-owned scratch (torch caching allocator => no cudaMalloc on the hot path)."""
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
