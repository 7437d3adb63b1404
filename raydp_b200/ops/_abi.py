"""ctypes signatures of ``include/raydp_b200.h`` (kept in one place, checked at load time)."""
import ctypes as C

VP, I, I64, F = C.c_void_p, C.c_int, C.c_int64, C.c_float

SIGNATURES = {
    "rdp_last_error": (C.c_char_p, []),
    "rdp_cuda_init": (I, [I]),
    "rdp_num_sms": (I, []),
    "rdp_gemm_tn": (I, [VP, I, VP, I, VP, I, I, I, I, I, VP, I, VP, I, I, I, VP, VP]),
    "rdp_gemm_wgrad": (I, [VP, I, VP, I, VP, I, I, I, I, VP]),
    "rdp_gemm_set_dynamic_tiles": (I, [I]),
    "rdp_gemm2_tn": (I, [VP, I, VP, I, VP, I, I, I, I, VP, I, I, VP]),
    "rdp_gemm2_wgrad": (I, [VP, I, VP, I, VP, I, I, I, I, I, VP]),
    "rdp_bn_fold": (I, [VP, F, VP, VP, VP, VP, F, F, I, VP, VP, VP, VP, VP, VP, I, I, VP, I, VP, I, VP, VP]),
    "rdp_cast_weight": (I, [VP, I, I, VP, I, VP, I, VP]),
    "rdp_bn_relu_bwd": (I, [VP, VP, VP, I, I, I, VP, VP, VP, VP, VP, VP, VP, I, VP]),
    "rdp_grad_fixup": (I, [VP, VP, VP, VP, I, I, VP]),
    "rdp_mlp_head": (I, [VP, I, I, I, VP, VP, VP, VP, VP, I, F, VP, VP, VP, VP, VP, VP, VP]),
    "rdp_adam_step": (I, [VP, VP, VP, VP, I64, F, F, F, F, F, I, F, I, VP, VP, VP, VP]),
    "rdp_sgd_step": (I, [VP, VP, VP, I64, F, F, F, F, I, I, F, VP, VP, VP, VP]),
    "rdp_relu_bwd_bias": (I, [VP, VP, VP, VP, I, I, I, VP]),
    "rdp_gemm2_nn": (I, [VP, I, VP, I, VP, I, I, I, I, I, VP]),
    "rdp_add_scalar": (I, [VP, F, VP]),
    "rdp_publish_host": (I, [VP, VP, I, VP]),
    "rdp_gemm_tn_bnbwd": (I, [VP, I, VP, I, VP, I, I, I, I, VP, I, VP, VP, VP, VP, VP]),
    "rdp_bn_bwd_coef": (I, [VP, I, VP, VP, I, I, I, VP, VP, VP, VP, F, VP, VP, VP, VP, VP, VP]),
    "rdp_memcpy_h2d_async": (I, [VP, VP, C.c_uint64, VP]),
    "rdp_memcpy_d2h_async": (I, [VP, VP, C.c_uint64, VP]),
    "rdp_ingest_pack": (I, [VP, VP, I, I64, I64, VP, I, I, VP, VP, VP, I, VP, VP]),
    "rdp_ingest_indices": (I, [VP, VP, I, I64, I64, VP, I, VP, VP, VP, VP, VP]),
    "rdp_ingest_pack_v2": (I, [VP, VP, VP, VP, VP, I, I64, I64, VP, I, I, VP, VP, F, VP, VP, I, I, VP, VP]),
    "rdp_ingest_indices_v2": (I, [VP, VP, VP, VP, I, I64, I64, VP, I, VP, VP, VP, VP, VP]),
    "rdp_ingest_stage": (I, [VP, VP, VP, I, VP]),
}

class OptArgs(C.Structure):
    _fields_ = [("kind", I), ("p", VP), ("m", VP), ("v", VP), ("lr", F), ("beta1", F), ("beta2", F),
                ("eps", F), ("weight_decay", F), ("momentum", F), ("dampening", F), ("adamw", I),
                ("nesterov", I), ("lr_ptr", VP), ("step_ptr", VP), ("shadow", VP)]


SIGNATURES.update({
    "rdp_comm_alloc": (I, [C.c_uint64, C.POINTER(VP)]),
    "rdp_comm_free": (I, [VP]),
    "rdp_comm_ipc_handle": (I, [VP, VP]),
    "rdp_comm_ipc_open": (I, [VP, C.POINTER(VP)]),
    "rdp_comm_ipc_close": (I, [VP]),
    "rdp_comm_can_access_peer": (I, [I, I]),
    "rdp_allreduce": (I, [C.POINTER(VP), C.c_uint64, I, I, I, C.POINTER(VP), C.POINTER(VP), C.POINTER(VP),
                          C.POINTER(VP), C.POINTER(OptArgs), I64, F, I, VP]),
})

SIGNATURES.update({
    "rdp_allreduce_ll": (I, [C.POINTER(VP), C.c_uint64, I, I, I, C.POINTER(VP), C.POINTER(VP), C.POINTER(VP),
                             C.POINTER(VP), C.POINTER(OptArgs), I64, F, VP]),
    "rdp_allreduce_bf16": (I, [C.POINTER(VP), C.c_uint64, I, I, I, C.POINTER(VP), C.POINTER(VP), C.POINTER(VP),
                               C.POINTER(VP), C.POINTER(OptArgs), I64, F, VP]),
    "rdp_comm_set_timeout": (I, [C.c_double, VP]),
    "rdp_comm_barrier": (I, [C.POINTER(VP), C.c_uint64, I, I, VP, VP]),
    "rdp_alltoall": (I, [VP, C.POINTER(VP), C.POINTER(VP), I, I, I64, I64, VP, VP, VP]),
    "rdp_allreduce_nvls": (I, [VP, VP, C.POINTER(VP), I, I, VP, VP, C.POINTER(OptArgs), VP, I64, F, I, VP]),
    "rdp_allreduce_nvls_bf16": (I, [VP, VP, C.POINTER(VP), I, I, VP, VP, C.POINTER(OptArgs), VP, VP, I64, F, I, VP]),
    "rdp_embedding_fwd": (I, [C.POINTER(VP), C.POINTER(I), C.POINTER(I), I, C.POINTER(VP), C.POINTER(VP),
                              C.POINTER(I), I, I, I, I, VP]),
    "rdp_embedding_bwd_sgd": (I, [C.POINTER(VP), C.POINTER(I), C.POINTER(I), I, C.POINTER(VP), C.POINTER(VP),
                                  C.POINTER(I), I, I, I, I, F, F, VP, VP]),
    "rdp_embedding_bag_fwd": (I, [C.POINTER(VP), C.POINTER(I), C.POINTER(I), C.POINTER(I), I, C.POINTER(VP),
                                  C.POINTER(VP), C.POINTER(I), C.POINTER(VP), I, I, I, I, I, VP]),
    "rdp_embedding_bag_bwd": (I, [C.POINTER(VP), C.POINTER(I), C.POINTER(I), C.POINTER(I), I, C.POINTER(VP),
                                  C.POINTER(VP), C.POINTER(I), C.POINTER(VP), I, I, I, I, I, F,
                                  VP, C.c_longlong, C.POINTER(VP), VP, VP, C.POINTER(VP), C.POINTER(I64),
                                  I, F, F, F, F, F, VP, VP, VP]),
    "rdp_gemm_wgrad_det": (I, [VP, I, VP, I, VP, I, I, I, I, VP, C.c_longlong, VP]),
    "rdp_gemm_wgrad_partial": (I, [VP, I, VP, I, VP, I, I, I, I, VP, C.c_longlong, C.POINTER(I), VP]),
    "rdp_wgrad_reduce_many": (I, [C.POINTER(VP), C.POINTER(C.c_longlong), C.POINTER(I), C.POINTER(VP), I, VP]),
    "rdp_colsum_ordered": (I, [C.POINTER(VP), C.POINTER(I), C.POINTER(I), C.POINTER(VP), I, I, VP, C.c_longlong, VP, I,
                               VP]),
    "rdp_cast_weight_many": (I, [C.POINTER(VP), C.POINTER(I), C.POINTER(I), C.POINTER(VP), C.POINTER(I), C.POINTER(VP),
                                 C.POINTER(I), C.POINTER(VP), C.POINTER(VP), I, VP]),
    "rdp_gbt_hist": (I, [VP, VP, VP, VP, VP, C.c_longlong, I, I, I, VP]),
    "rdp_gbt_partition": (I, [VP, VP, VP, VP, VP, VP, VP, VP, VP, VP, C.c_longlong, I, VP]),
    "rdp_dot_interaction_fwd": (I, [VP, VP, I, I, I, I, VP]),
    "rdp_dot_interaction_bwd": (I, [VP, VP, VP, I, I, I, I, VP]),
    "rdp_bce_logits": (I, [VP, I, VP, VP, I, I, VP, I, F, VP, VP]),
})

SIGNATURES.update({
    "rdp_symm_supported": (I, [I]),
    "rdp_symm_granularity": (I, [I, I, I, C.POINTER(C.c_uint64)]),
    "rdp_symm_alloc": (I, [C.c_uint64, I, C.POINTER(VP), C.POINTER(VP)]),
    "rdp_symm_export_fd": (I, [VP, C.POINTER(I)]),
    "rdp_symm_import_fd": (I, [I, C.c_uint64, I, C.POINTER(VP), C.POINTER(VP)]),
    "rdp_symm_free": (I, [VP]),
    "rdp_mc_create": (I, [C.c_uint64, I, C.POINTER(VP)]),
    "rdp_mc_import_fd": (I, [I, C.c_uint64, C.POINTER(VP)]),
    "rdp_mc_add_device": (I, [VP, I]),
    "rdp_mc_bind": (I, [VP, VP]),
    "rdp_mc_map": (I, [VP, I, C.POINTER(VP)]),
    "rdp_close_fd": (I, [I]),
})

OPTIONAL = {}


def declare(lib) -> None:
    missing = []
    for name, (res, args) in {**SIGNATURES, **OPTIONAL}.items():
        fn = getattr(lib, name, None)
        if fn is None:
            if name in SIGNATURES:
                missing.append(name)
            continue
        fn.restype = res
        fn.argtypes = args
    if missing:
        raise RuntimeError(f"librdp_kernels.so is stale, missing symbols: {missing}; rebuild it")
