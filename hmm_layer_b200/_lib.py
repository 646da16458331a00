"""ctypes binding of libhmm_b200.so (the C ABI declared in include/hmm_b200.h).

There is no CPU or PyTorch fallback: if the shared library is missing the import of any hot-path function fails
loudly, and every entry point requires CUDA tensors."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HMM_B200_LIB", os.path.join(_HERE, "libhmm_b200.so"))   # override: build-variant experiments

EXPORTS = [
    "hmm_b200_abi_version", "hmm_b200_last_error", "hmm_b200_set_tuning", "hmm_b200_set_fused_mode", "hmm_b200_profile_enable", "hmm_b200_profile_read", "hmm_b200_debug_phases",
    "hmm_b200_genepred_workspace_size", "hmm_b200_genepred_prepare_codon", "hmm_b200_genepred_run",
    "hmm_b200_genepred_host_scratch_size", "hmm_b200_genepred_run_host",
    "hmm_b200_genepred_run_compact", "hmm_b200_genepred_run_host_compact",
    "hmm_b200_genepred_grad_workspace_size", "hmm_b200_genepred_grad",
    "hmm_b200_genepred_viterbi_workspace_size", "hmm_b200_genepred_viterbi",
    "hmm_b200_genepred_params", "hmm_b200_genepred_params_grad",
    "hmm_b200_genepred_emission", "hmm_b200_emission_general", "hmm_b200_genepred_shard_reduce", "hmm_b200_genepred_shard_finalize",
    "hmm_b200_genepred_shard_reduce_x", "hmm_b200_genepred_shard_finalize_x",
    "hmm_b200_genepred_shard_fused_chunk", "hmm_b200_genepred_shard_run_x", "hmm_b200_genepred_run_noalloc",
    "hmm_b200_dense_workspace_size", "hmm_b200_dense_run", "hmm_b200_dense_workspace_size_onehot", "hmm_b200_dense_run_onehot", "hmm_b200_cell_step",
]

_lib = None


class HmmB200Error(RuntimeError):
    pass


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `bash hmm_layer_b200/csrc/build.sh` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    c_fp = ctypes.c_void_p
    lib.hmm_b200_abi_version.restype = ctypes.c_int
    lib.hmm_b200_last_error.restype = ctypes.c_char_p
    lib.hmm_b200_set_tuning.argtypes = [ctypes.c_int] * 3
    lib.hmm_b200_set_fused_mode.argtypes = [ctypes.c_int]
    lib.hmm_b200_profile_enable.argtypes = [ctypes.c_int]
    lib.hmm_b200_debug_phases.argtypes = [ctypes.c_void_p]
    lib.hmm_b200_profile_read.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.c_int]
    lib.hmm_b200_genepred_workspace_size.restype = ctypes.c_size_t
    lib.hmm_b200_genepred_workspace_size.argtypes = [ctypes.c_int] * 3
    lib.hmm_b200_genepred_prepare_codon.argtypes = [c_fp, c_fp, c_fp]
    lib.hmm_b200_genepred_run.argtypes = [c_fp, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_fp, c_fp, c_fp, c_fp, c_fp,
                                          ctypes.c_uint, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_size_t, c_fp]
    lib.hmm_b200_genepred_run_noalloc.argtypes = [c_fp, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_fp, c_fp, c_fp, c_fp, c_fp,
                                                  ctypes.c_uint, c_fp, c_fp, c_fp, c_fp, c_fp]
    lib.hmm_b200_genepred_host_scratch_size.restype = ctypes.c_size_t
    lib.hmm_b200_genepred_host_scratch_size.argtypes = [ctypes.c_int] * 3
    lib.hmm_b200_genepred_run_host.argtypes = lib.hmm_b200_genepred_run.argtypes
    lib.hmm_b200_genepred_run_compact.argtypes = [c_fp] + lib.hmm_b200_genepred_run.argtypes
    lib.hmm_b200_genepred_run_host_compact.argtypes = [c_fp] + lib.hmm_b200_genepred_run.argtypes
    lib.hmm_b200_genepred_emission.argtypes = [c_fp, ctypes.c_int, ctypes.c_int, ctypes.c_int, c_fp, c_fp, c_fp,
                                               ctypes.c_uint, c_fp, c_fp]
    ci = ctypes.c_int
    lib.hmm_b200_emission_general.argtypes = [c_fp, ci, ci, ci, ci, ci, ci, ci, c_fp, c_fp, c_fp, ctypes.c_uint, c_fp, c_fp]
    lib.hmm_b200_genepred_params.argtypes = [c_fp, c_fp, c_fp, ci, c_fp, c_fp, c_fp, c_fp]
    lib.hmm_b200_genepred_params_grad.argtypes = [c_fp, c_fp, c_fp, ci, c_fp, c_fp, c_fp, c_fp, c_fp, c_fp, c_fp]
    lib.hmm_b200_genepred_viterbi_workspace_size.restype = ctypes.c_size_t
    lib.hmm_b200_genepred_viterbi_workspace_size.argtypes = [ctypes.c_int] * 3
    lib.hmm_b200_genepred_viterbi.argtypes = [c_fp, ci, ci, ci, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_uint, c_fp, c_fp,
                                              c_fp, ctypes.c_size_t, c_fp]
    lib.hmm_b200_genepred_grad_workspace_size.restype = ctypes.c_size_t
    lib.hmm_b200_genepred_grad_workspace_size.argtypes = [ctypes.c_int] * 3
    lib.hmm_b200_genepred_grad.argtypes = [c_fp, ci, ci, ci, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_uint, c_fp, c_fp, c_fp,
                                           c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_size_t, c_fp]
    lib.hmm_b200_genepred_shard_reduce.argtypes = [c_fp, ci, ci, ci, ci, ci, c_fp, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_uint,
                                                   c_fp, c_fp, ctypes.c_size_t, c_fp]
    lib.hmm_b200_genepred_shard_finalize.argtypes = [c_fp, ci, ci, ci, ci, ci, c_fp, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_uint,
                                                     c_fp, ci, ci, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_size_t, c_fp]
    lib.hmm_b200_genepred_shard_reduce_x.argtypes = [c_fp, ci, ci, ci, ci, ci, c_fp, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_uint,
                                                     c_fp, c_fp, c_fp, ci, ci, ctypes.c_uint, c_fp, ctypes.c_size_t, c_fp]
    lib.hmm_b200_genepred_shard_finalize_x.argtypes = [c_fp, ci, ci, ci, ci, ci, c_fp, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_uint,
                                                       c_fp, c_fp, ci, ci, ctypes.c_uint, c_fp, c_fp, c_fp, c_fp, c_fp,
                                                       ctypes.c_size_t, c_fp]
    lib.hmm_b200_genepred_shard_fused_chunk.argtypes = [ci, ci, ci, ci]
    lib.hmm_b200_genepred_shard_run_x.argtypes = [c_fp, ci, ci, ci, ci, ci, c_fp, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_uint,
                                                  c_fp, c_fp, ci, ci, ctypes.c_uint, c_fp, c_fp, c_fp, ctypes.c_size_t, c_fp]
    lib.hmm_b200_dense_workspace_size.restype = ctypes.c_size_t
    lib.hmm_b200_dense_workspace_size.argtypes = [ctypes.c_int] * 4
    lib.hmm_b200_dense_run.argtypes = [c_fp, c_fp, c_fp, c_fp, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_uint, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_size_t, c_fp]
    lib.hmm_b200_dense_workspace_size_onehot.restype = ctypes.c_size_t
    lib.hmm_b200_dense_workspace_size_onehot.argtypes = [ctypes.c_int] * 4
    lib.hmm_b200_dense_run_onehot.argtypes = [c_fp, c_fp, ctypes.c_int, c_fp, c_fp, c_fp, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                              ctypes.c_int, ctypes.c_uint, c_fp, c_fp, c_fp, c_fp, c_fp, ctypes.c_size_t, c_fp]
    lib.hmm_b200_cell_step.argtypes = [c_fp, c_fp, c_fp, c_fp] + [ctypes.c_int] * 6 + [c_fp, c_fp, c_fp, c_fp]
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().hmm_b200_last_error().decode("utf-8", "replace")
        raise HmmB200Error(f"libhmm_b200 error {rc}: {msg}")
