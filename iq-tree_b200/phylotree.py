"""ctypes binding of the host PhyloTree mirror (include/lktree.h) and of the engine C ABI (include/lkcuda.h).

The method names follow the reference's PhyloTree likelihood API (phylotree.h:307 ff.): computeLikelihood,
computeLikelihoodBranch, computeLikelihoodDerv, computeLikelihoodFromBuffer, optimizeOneBranch,
optimizeAllBranches, clearAllPartialLH, initializeAllPartialLh, computePatternLhCat.  All arithmetic happens
behind the C ABI on the GPU; this file only marshals numpy arrays.  There is no CPU fallback: the product
library is the CUDA one, and loading it or creating an engine without a device raises.
"""
import ctypes as C
import os

import numpy as np

from . import _build

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)
c_int16_p = C.POINTER(C.c_int16)
c_uint8_p = C.POINTER(C.c_uint8)
c_int64_p = C.POINTER(C.c_int64)
c_uint64_p = C.POINTER(C.c_uint64)

ALLREDUCE_FN = C.CFUNCTYPE(None, C.c_void_p, c_double_p, C.c_void_p, C.c_int)

NRESULT = 8
R_LNL, R_DF, R_DDF, R_PROB_CONST, R_NREPAIRED, R_SCALE_SUM = 0, 1, 2, 3, 4, 5
KIND_PREP, KIND_PARTIAL, KIND_BRANCH, KIND_THETA, KIND_DERV, KIND_FROMBUF, KIND_NEWTON, KIND_MISC = range(8)
KIND_NAMES = ["prep", "partial", "branch", "theta", "derv", "frombuf", "newton", "misc", "rell", "em"]
FAMILY_PLAIN, FAMILY_MIXRATE, FAMILY_MIXTURE = 0, 1, 2
SEQ_DNA, SEQ_PROTEIN, SEQ_BINARY, SEQ_MORPH, SEQ_MULTISTATE, SEQ_CODON = range(6)
WSL_RATECAT, WSL_MIXTURE, WSL_MIXTURE_RATECAT = 2, 3, 4

# name -> (restype, argtypes); every symbol include/lkcuda.h and include/lktree.h declare
LKCUDA_SYMBOLS = {
    "lkcuda_device_count": (C.c_int, [c_int32_p]),
    "lkcuda_create": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "lkcuda_destroy": (C.c_int, [C.c_void_p]),
    "lkcuda_last_error": (C.c_char_p, [C.c_void_p]),
    "lkcuda_memory_bytes": (C.c_int, [C.c_void_p, c_uint64_p]),
    "lkcuda_device_malloc": (C.c_int, [C.c_int32, C.c_uint64, C.POINTER(C.c_void_p)]),
    "lkcuda_device_free": (C.c_int, [C.c_int32, C.c_void_p]),
    "lkcuda_host_register": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "lkcuda_host_unregister": (C.c_int, [C.c_void_p, C.c_void_p]),
    "lkcuda_arena_layout": (C.c_int, [C.c_void_p, c_int64_p, c_int64_p]),
    "lkcuda_attach_arena": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int32]),
    "lkcuda_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "lkcuda_synchronize": (C.c_int, [C.c_void_p]),
    "lkcuda_set_tip_states": (C.c_int, [C.c_void_p, c_uint8_p]),
    "lkcuda_set_tip_states_async": (C.c_int, [C.c_void_p, c_uint8_p, C.c_int32]),
    "lkcuda_set_pattern_weights": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "lkcuda_set_model": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p, c_int32_p, c_double_p, c_double_p, c_double_p]),
    "lkcuda_partial_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, c_int32_p, C.c_int32]),
    "lkcuda_theta_newton_async": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_int32, C.c_double, C.c_double, C.c_int32, C.c_int32]),
    "lkcuda_lnl_from_buffer_slot": (C.c_int, [C.c_void_p, C.c_int32, c_double_p]),
    "lkcuda_len_slots_set": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, c_double_p]),
    "lkcuda_len_slots_fetch": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "lkcuda_partial_batch_lnl": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, c_int32_p, C.c_int32, C.c_int32, C.c_int32, C.c_double, c_double_p]),
    "lkcuda_partial_multi": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, c_int32_p, c_double_p]),
    "lkcuda_branch_lnl": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_double, c_double_p]),
    "lkcuda_pattern_lh_cat": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_double]),
    "lkcuda_theta_build": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "lkcuda_theta_derv": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_double, c_double_p]),
    "lkcuda_derv": (C.c_int, [C.c_void_p, C.c_double, c_double_p]),
    "lkcuda_derv_noalloc": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, c_double_p, c_double_p, c_double_p,
                            C.c_double, C.c_void_p]),
    "lkcuda_lnl_from_buffer": (C.c_int, [C.c_void_p, C.c_double, c_double_p]),
    "lkcuda_newton_branch": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32, c_double_p]),
    "lkcuda_theta_newton": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32, c_double_p]),
    "lkcuda_read_result": (C.c_int, [C.c_void_p, c_double_p]),
    "lkcuda_wait_result": (C.c_int, [C.c_void_p, c_double_p]),
    "lkcuda_peer_handle": (C.c_int, [C.c_void_p, c_uint8_p, C.POINTER(C.c_void_p)]),
    "lkcuda_attach_peers": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, c_uint8_p, C.POINTER(C.c_void_p)]),
    "lkcuda_detach_peers": (C.c_int, [C.c_void_p]),
    "lkcuda_result_device_ptr": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "lkcuda_set_site_models": (C.c_int, [C.c_void_p, c_int32_p, C.POINTER(C.c_uint64)]),
    "lkcuda_set_boot_samples": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.c_int32, C.c_int64]),
    "lkcuda_boot_storage": (C.c_int, [C.c_void_p, c_int32_p, c_uint64_p]),
    "lkcuda_rell": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, c_double_p]),
    "lkcuda_rell_fetch": (C.c_int, [C.c_void_p, c_double_p]),
    "lkcuda_rell_device_ptr": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    "lkcuda_lh_cat_collapse": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, c_int32_p]),
    "lkcuda_lh_cat_posterior_weights": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "lkcuda_lh_cat_em": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "lkcuda_lh_cat_em_ex": (C.c_int, [C.c_void_p, c_double_p, C.c_int32, c_double_p]),
    "lkcuda_lh_cat_pattern_rates": (C.c_int, [C.c_void_p, c_double_p, C.c_int32, c_double_p, c_int32_p]),
    "lkcuda_fetch_lh_cat_collapsed": (C.c_int, [C.c_void_p, c_double_p]),
    "lkcuda_set_result_buffer": (C.c_int, [C.c_void_p, C.c_void_p]),
    "lkcuda_fetch_pattern_lh": (C.c_int, [C.c_void_p, c_double_p]),
    "lkcuda_fetch_pattern_lh_cat": (C.c_int, [C.c_void_p, c_double_p]),
    "lkcuda_fetch_pattern_lh_corrected_async": (C.c_int, [C.c_void_p, C.c_int32, c_double_p]),
    "lkcuda_fetch_wait": (C.c_int, [C.c_void_p]),
    "lkcuda_fetch_pattern_lh_corrected": (C.c_int, [C.c_void_p, C.c_int32, c_double_p]),
    "lkcuda_fetch_vector": (C.c_int, [C.c_void_p, C.c_int32, c_double_p, c_int16_p]),
    "lkcuda_store_vector": (C.c_int, [C.c_void_p, C.c_int32, c_double_p, c_int16_p]),
    "lkcuda_fetch_theta": (C.c_int, [C.c_void_p, c_double_p]),
    "lkcuda_vector_scale_factor": (C.c_int, [C.c_void_p, C.c_int32, c_double_p]),
    "lkcuda_vector_device_ptr": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "lkcuda_launch_count": (C.c_int, [C.c_void_p, c_uint64_p]),
    "lkcuda_algorithmic_bytes": (C.c_int, [C.c_void_p, c_uint64_p]),
    "lkcuda_reset_counters": (C.c_int, [C.c_void_p]),
    "lkcuda_set_profiling": (C.c_int, [C.c_void_p, C.c_int32]),
    "lkcuda_kernel_stats": (C.c_int, [C.c_void_p, C.c_int32, c_uint64_p, c_uint64_p, c_double_p]),
    "lkcuda_kernel_flops": (C.c_int, [C.c_void_p, C.c_int32, c_uint64_p]),
}

LKTREE_SYMBOLS = {
    "lktree_create": (C.c_void_p, []),
    "lktree_destroy": (None, [C.c_void_p]),
    "lktree_last_error": (C.c_char_p, [C.c_void_p]),
    "lktree_set_alignment": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, c_uint8_p, c_double_p, C.POINTER(C.c_char_p), C.c_int32, C.c_int32, C.c_int32]),
    "lktree_set_const_char": (C.c_int, [C.c_void_p, c_int32_p]),
    "lktree_set_unobserved": (C.c_int, [C.c_void_p, c_uint8_p, C.c_int32]),
    "lktree_set_model": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, C.c_double]),
    "lktree_set_site_models": (C.c_int, [C.c_void_p, c_int32_p]),
    "lktree_read_newick": (C.c_int, [C.c_void_p, C.c_char_p]),
    "lktree_read_adjacency": (C.c_int, [C.c_void_p, C.c_char_p]),
    "lktree_get_tree_string": (C.c_int, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "lktree_set_devices": (C.c_int, [C.c_void_p, c_int32_p, C.c_int32]),
    "lktree_set_process_shard": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    "lktree_set_allreduce": (C.c_int, [C.c_void_p, ALLREDUCE_FN, C.c_void_p]),
    "lktree_set_boot_samples": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.c_int32]),
    "lktree_compute_rell": (C.c_int, [C.c_void_p, c_double_p]),
    "lktree_optimize_mixture_weights": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "lktree_compute_pattern_rates": (C.c_int, [C.c_void_p, c_double_p, c_int32_p]),
    "lktree_peer_handle": (C.c_int, [C.c_void_p, c_uint8_p]),
    "lktree_attach_peers": (C.c_int, [C.c_void_p, c_uint8_p]),
    "lktree_detach_peers": (C.c_int, [C.c_void_p]),
    "lktree_peers_attached": (C.c_int, [C.c_void_p]),
    "lktree_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "lktree_set_params": (C.c_int, [C.c_void_p, C.c_double, C.c_double, C.c_int32, C.c_int32]),
    "lktree_initialize_all_partial_lh": (C.c_int, [C.c_void_p]),
    "lktree_delete_all_partial_lh": (C.c_int, [C.c_void_p]),
    "lktree_clear_all_partial_lh": (C.c_int, [C.c_void_p]),
    "lktree_compute_likelihood": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "lktree_compute_likelihood_async_fetch": (C.c_int, [C.c_void_p, c_double_p, c_double_p]),
    "lktree_wait_pattern_lh": (C.c_int, [C.c_void_p]),
    "lktree_compute_likelihood_branch": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, c_double_p]),
    "lktree_compute_likelihood_derv": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, c_double_p, c_double_p]),
    "lktree_nni_queued_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "lktree_bench_steps": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, c_double_p]),
    "lktree_compute_likelihood_from_buffer": (C.c_int, [C.c_void_p, c_double_p]),
    "lktree_optimize_one_branch": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, c_double_p]),
    "lktree_get_best_nni_for_bran": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, c_double_p]),
    "lktree_optimize_all_branches": (C.c_int, [C.c_void_p, C.c_int32, C.c_double, C.c_int32, c_double_p]),
    "lktree_set_branch_length": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_double]),
    "lktree_get_branch_length": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, c_double_p]),
    "lktree_get_traversal": (C.c_int, [C.c_void_p, c_int32_p, C.c_int32, c_int32_p]),
    "lktree_branch_vector": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, c_int32_p, c_int32_p]),
    "lktree_root_branch": (C.c_int, [C.c_void_p, c_int32_p, c_int32_p]),
    "lktree_fetch_tip_partial_lh": (C.c_int, [C.c_void_p, c_double_p, C.c_int64, c_int64_p]),
    "lktree_fetch_ptn_invar": (C.c_int, [C.c_void_p, c_double_p]),
    "lktree_compute_pattern_lh_cat": (C.c_int, [C.c_void_p, C.c_int32, c_double_p, c_int32_p]),
    "lktree_engine": (C.c_void_p, [C.c_void_p, C.c_int32]),
    "lktree_local_patterns": (C.c_int, [C.c_void_p, c_int64_p, c_int64_p]),
    "lktree_shard_patterns": (C.c_int, [C.c_void_p, C.c_int32, c_int64_p, c_int64_p]),
    "lktree_num_partial_lh_computations": (C.c_uint64, [C.c_void_p]),
    "lktree_optimize_rate_free_em": (C.c_int, [C.c_void_p, C.c_int32, c_double_p, c_double_p, c_double_p]),
    "lktree_set_parsimony_sites": (C.c_int, [C.c_void_p, C.c_int64, c_uint8_p, C.c_int64, C.c_int32]),
    "lktree_compute_parsimony": (C.c_int, [C.c_void_p, c_int32_p]),
    "lktree_compute_parsimony_branch": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, c_int32_p, c_int32_p]),
    "lktree_fix_negative_branch": (C.c_int, [C.c_void_p, C.c_int32, c_int32_p]),
    "lktree_compute_parsimony_tree": (C.c_int, [C.c_void_p, c_int32_p, c_int32_p]),
    "lktree_get_branches": (C.c_int, [C.c_void_p, c_int32_p, C.c_int32, c_int32_p]),
    "lktree_get_adjacency": (C.c_int, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "lktree_pars_engine": (C.c_void_p, [C.c_void_p]),
}

# include/lkpars.h
LKPARS_SYMBOLS = {
    "lkpars_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int64, c_uint8_p, c_uint64_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "lkpars_destroy": (C.c_int, [C.c_void_p]),
    "lkpars_last_error": (C.c_char_p, [C.c_void_p]),
    "lkpars_run": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.POINTER(C.c_uint32)]),
    "lkpars_words": (C.c_int, [C.c_void_p, c_int64_p]),
    "lkpars_fetch_vector": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_uint32)]),
    "lkpars_launch_count": (C.c_int, [C.c_void_p, c_uint64_p, c_double_p]),
    "lkpars_set_profiling": (C.c_int, [C.c_void_p, C.c_int32]),
}


class LkError(RuntimeError):
    pass


def _bind(lib):
    for table in (LKCUDA_SYMBOLS, LKTREE_SYMBOLS, LKPARS_SYMBOLS):
        for name, (res, args) in table.items():
            fn = getattr(lib, name)      # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
    return lib


_libs = {}


def load_library(path=None):
    """Load (building first if sources are newer) the CUDA product library through the harness library that drives it
    (harness/: the PhyloTree mirror used by tests and bench.py; it links iq-tree_b200/lib/liblktree_cuda.so).  Never falls back to a
    CPU library."""
    if path is None:
        path = _build.build_harness()
    path = os.path.abspath(path)
    if path not in _libs:
        if not os.path.exists(path):
            raise LkError("native library %s is missing: run __graft_entry__.build()" % path)
        _libs[path] = _bind(C.CDLL(path))
    return _libs[path]


def _dp(a):
    return a.ctypes.data_as(c_double_p)


def _f64(a, n=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if n is not None and a.size != n:
        raise ValueError("expected %d doubles, got %d" % (n, a.size))
    return a


class PhyloTree:
    """Host-side handle with the reference's PhyloTree likelihood API, backed by the CUDA engine."""

    def __init__(self, lib=None):
        self.lib = lib if lib is not None else load_library()
        self.h = self.lib.lktree_create()
        if not self.h:
            raise LkError("lktree_create failed")
        self._cb = None
        self.ntaxa = 0
        self.nptn = 0
        self.nstates = 0
        self.block = 0
        self.nblocks = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.lktree_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise LkError(self.lib.lktree_last_error(self.h).decode())

    # ---- inputs -------------------------------------------------------------------------------------------
    def setAlignment(self, states, ptn_freq, names=None, seq_type=SEQ_DNA, nstates=4, state_unknown=18):
        states = np.ascontiguousarray(states, dtype=np.uint8)
        ntaxa, nptn = states.shape
        freq = _f64(ptn_freq, nptn)
        arr = None
        if names is not None:
            arr = (C.c_char_p * ntaxa)(*[n.encode() for n in names])
        self._ck(self.lib.lktree_set_alignment(self.h, ntaxa, nptn, states.ctypes.data_as(c_uint8_p), _dp(freq), arr,
                                               seq_type, nstates, state_unknown))
        self.ntaxa, self.nptn, self.nstates = ntaxa, nptn, nstates
        self.nunobs = 0

    def setConstChar(self, const_char):
        cc = np.ascontiguousarray(const_char, dtype=np.int32)
        self._ck(self.lib.lktree_set_const_char(self.h, cc.ctypes.data_as(c_int32_p)))

    def setUnobservedPatterns(self, unobs):
        u = np.ascontiguousarray(unobs, dtype=np.uint8)
        self._ck(self.lib.lktree_set_unobserved(self.h, u.ctypes.data_as(c_uint8_p), u.size))
        self.nunobs = int(u.size)

    def setModel(self, family, ncat, nmix, eval_, evec, inv_evec, rates, props, mixprops=None, state_freq=None, p_invar=0.0):
        S = self.nstates
        ev, evc, inv = _f64(eval_, nmix * S), _f64(evec, nmix * S * S), _f64(inv_evec, nmix * S * S)
        r, p = _f64(rates, ncat), _f64(props, ncat)
        mp = _f64(mixprops, nmix) if mixprops is not None else None
        sf = _f64(state_freq, S) if state_freq is not None else None
        self._ck(self.lib.lktree_set_model(self.h, family, ncat, nmix, _dp(ev), _dp(evc), _dp(inv), _dp(r), _dp(p),
                                           _dp(mp) if mp is not None else None, _dp(sf) if sf is not None else None, float(p_invar)))
        self.family, self.ncat, self.nmix = family, ncat, nmix
        self.nblocks = ncat * nmix if family == FAMILY_MIXTURE else ncat
        self.block = self.nblocks * S

    def readTreeString(self, newick):
        self._ck(self.lib.lktree_read_newick(self.h, newick.encode()))

    def readAdjacency(self, text):
        self._ck(self.lib.lktree_read_adjacency(self.h, text.encode()))

    def getTreeString(self):
        buf = C.create_string_buffer(1 << 22)
        self._ck(self.lib.lktree_get_tree_string(self.h, buf, len(buf)))
        return buf.value.decode()

    # ---- placement ----------------------------------------------------------------------------------------
    def setDevices(self, devices):
        d = np.ascontiguousarray(devices, dtype=np.int32)
        self._ck(self.lib.lktree_set_devices(self.h, d.ctypes.data_as(c_int32_p), d.size))

    def setProcessShard(self, rank, world):
        self._ck(self.lib.lktree_set_process_shard(self.h, rank, world))

    def setAllReduce(self, fn):
        """fn(host_array_or_None, dev_ptr_or_None, n): sum n doubles over the ranks in place."""
        if fn is None:
            self._cb = None
            self._ck(self.lib.lktree_set_allreduce(self.h, C.cast(None, ALLREDUCE_FN), None))
            return

        def tramp(user, host_buf, dev_buf, n):
            if host_buf:
                fn(np.ctypeslib.as_array(host_buf, shape=(n,)), None, n)
            else:
                fn(None, dev_buf, n)

        self._cb = ALLREDUCE_FN(tramp)
        self._ck(self.lib.lktree_set_allreduce(self.h, self._cb, None))

    # ---- RELL and the _pattern_lh_cat consumers --------------------------------------------------
    def setBootSamples(self, samples):
        """samples: float32[nsamples][nptn] bootstrap pattern frequencies (IQTree::boot_samples, iqtree.cpp:300-340)"""
        a = np.ascontiguousarray(samples, dtype=np.float32)
        self._nboot = a.shape[0]
        self._ck(self.lib.lktree_set_boot_samples(self.h, a.ctypes.data_as(C.POINTER(C.c_float)), a.shape[0]))

    def computeRell(self):
        out = np.zeros(self._nboot, dtype=np.float64)
        self._ck(self.lib.lktree_compute_rell(self.h, out.ctypes.data_as(c_double_p)))
        return out

    def optimizeMixtureWeights(self, nmix):
        """ModelMixture::optimizeWeights: (lnL after the EM, new class weights)"""
        lnl = C.c_double()
        w = np.zeros(nmix, dtype=np.float64)
        self._ck(self.lib.lktree_optimize_mixture_weights(self.h, C.byref(lnl), w.ctypes.data_as(c_double_p)))
        return lnl.value, w

    def computePatternRates(self):
        b0, b1 = self.localPatterns()
        r = np.zeros(b1 - b0, dtype=np.float64)
        c = np.zeros(b1 - b0, dtype=np.int32)
        self._ck(self.lib.lktree_compute_pattern_rates(self.h, r.ctypes.data_as(c_double_p), c.ctypes.data_as(c_int32_p)))
        return r, c

    def setSiteModels(self, ptn_model):
        """family 3 (site-specific frequency models): ptn_model[nptn] -> index of the pattern's eigensystem"""
        a = np.ascontiguousarray(ptn_model, dtype=np.int32)
        self._ck(self.lib.lktree_set_site_models(self.h, a.ctypes.data_as(c_int32_p)))

    def peerHandle(self):
        """CUDA IPC handle (64 bytes, numpy uint8) of this rank's all-reduce mailbox; exchange with every rank, then attachPeers()."""
        h = np.zeros(64, dtype=np.uint8)
        self._ck(self.lib.lktree_peer_handle(self.h, h.ctypes.data_as(c_uint8_p)))
        return h

    def attachPeers(self, handles):
        """handles: uint8[world][64] in rank order. From now on every reduction kernel returns the sums over all ranks."""
        a = np.ascontiguousarray(handles, dtype=np.uint8)
        self._ck(self.lib.lktree_attach_peers(self.h, a.ctypes.data_as(c_uint8_p)))

    def detachPeers(self):
        self._ck(self.lib.lktree_detach_peers(self.h))

    def peersAttached(self):
        return bool(self.lib.lktree_peers_attached(self.h))

    def setStream(self, cuda_stream):
        self._ck(self.lib.lktree_set_stream(self.h, C.c_void_p(cuda_stream)))

    def setParams(self, min_branch_length=1e-6, max_branch_length=100.0, device_newton=False, generic_kernel=False):
        self._ck(self.lib.lktree_set_params(self.h, min_branch_length, max_branch_length, int(device_newton), int(generic_kernel)))

    # ---- PhyloTree API ------------------------------------------------------------------------------------
    def initializeAllPartialLh(self):
        self._ck(self.lib.lktree_initialize_all_partial_lh(self.h))

    def deleteAllPartialLh(self):
        self._ck(self.lib.lktree_delete_all_partial_lh(self.h))

    def clearAllPartialLH(self):
        self._ck(self.lib.lktree_clear_all_partial_lh(self.h))

    def localPatterns(self):
        b, e = C.c_int64(), C.c_int64()
        self._ck(self.lib.lktree_local_patterns(self.h, C.byref(b), C.byref(e)))
        return b.value, e.value

    def computeLikelihood(self, pattern_lh=False):
        lnl = C.c_double()
        if pattern_lh:
            b, e = self.localPatterns()
            out = np.zeros(max(e - b, 1), dtype=np.float64)
            self._ck(self.lib.lktree_compute_likelihood(self.h, C.byref(lnl), _dp(out)))
            return lnl.value, out[: e - b]
        self._ck(self.lib.lktree_compute_likelihood(self.h, C.byref(lnl), None))
        return lnl.value

    def computeLikelihoodBranch(self, node1, node2):
        lnl = C.c_double()
        self._ck(self.lib.lktree_compute_likelihood_branch(self.h, node1, node2, C.byref(lnl)))
        return lnl.value

    def computeLikelihoodDerv(self, node1, node2, reset_theta=True):
        df, ddf = C.c_double(), C.c_double()
        self._ck(self.lib.lktree_compute_likelihood_derv(self.h, node1, node2, int(reset_theta), C.byref(df), C.byref(ddf)))
        return df.value, ddf.value

    def nniQueuedStats(self):
        """(calls of getBestNNIForBran that ran queued through device length slots, calls redone step by step)"""
        a, b = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.lktree_nni_queued_stats(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def benchSteps(self, node1, node2, steps):
        """`steps` x { clearAllPartialLH; computeLikelihood; computeLikelihoodDerv(node1 - node2, theta rebuilt) } driven from C++"""
        out = (C.c_double * 3)()
        self._ck(self.lib.lktree_bench_steps(self.h, node1, node2, steps, out))
        return out[0], out[1], out[2]

    def computeLikelihoodFromBuffer(self):
        lnl = C.c_double()
        self._ck(self.lib.lktree_compute_likelihood_from_buffer(self.h, C.byref(lnl)))
        return lnl.value

    def optimizeOneBranch(self, node1, node2, max_nr_step=100):
        ln = C.c_double()
        self._ck(self.lib.lktree_optimize_one_branch(self.h, node1, node2, max_nr_step, C.byref(ln)))
        return ln.value

    def getBestNNIForBran(self, node1, node2, nni5=True):
        """-> [(newloglh, newLen[5], node1_nei, node2_nei)] * 2 (PhyloTree::getBestNNIForBran, phylotree.cpp:2873-3066)"""
        out = np.zeros(16, dtype=np.float64)
        self._ck(self.lib.lktree_get_best_nni_for_bran(self.h, node1, node2, int(nni5), out.ctypes.data_as(c_double_p)))
        return [(out[8 * c], out[8 * c + 1: 8 * c + 6].copy(), int(out[8 * c + 6]), int(out[8 * c + 7])) for c in range(2)]

    def optimizeAllBranches(self, iterations=100, tolerance=0.001, max_nr_step=100):
        lnl = C.c_double()
        self._ck(self.lib.lktree_optimize_all_branches(self.h, iterations, tolerance, max_nr_step, C.byref(lnl)))
        return lnl.value

    def computePatternLhCat(self, wsl=WSL_MIXTURE_RATECAT):
        b, e = self.localPatterns()
        out = np.zeros(max((e - b) * max(self.nblocks, 1), 1), dtype=np.float64)
        ncol = C.c_int32()
        self._ck(self.lib.lktree_compute_pattern_lh_cat(self.h, wsl, _dp(out), C.byref(ncol)))
        return out[: (e - b) * ncol.value].reshape(e - b, ncol.value)

    # ---- inspection ---------------------------------------------------------------------------------------
    def setBranchLength(self, node1, node2, length):
        self._ck(self.lib.lktree_set_branch_length(self.h, node1, node2, float(length)))

    def getBranchLength(self, node1, node2):
        v = C.c_double()
        self._ck(self.lib.lktree_get_branch_length(self.h, node1, node2, C.byref(v)))
        return v.value

    def getTraversal(self):
        cap = 4 * max(self.ntaxa, 2)
        pairs = np.zeros(2 * cap, dtype=np.int32)
        cnt = C.c_int32()
        self._ck(self.lib.lktree_get_traversal(self.h, pairs.ctypes.data_as(c_int32_p), cap, C.byref(cnt)))
        return pairs[: 2 * cnt.value].reshape(-1, 2).copy()

    def optimizeRateFreeEM(self, ncat, copy_precision=6):
        """RateFree::optimizeWithEM: (lnL, rates[ncat], props[ncat])"""
        lnl = C.c_double()
        r, p = np.zeros(ncat), np.zeros(ncat)
        self._ck(self.lib.lktree_optimize_rate_free_em(self.h, copy_precision, C.byref(lnl), r.ctypes.data_as(c_double_p), p.ctypes.data_as(c_double_p)))
        return lnl.value, r, p

    # ---- Fitch parsimony ----
    def setParsimonySites(self, site_states, aln_nsite, device=-1):
        """site_states[ntaxa][nsites]: the informative sites in Alignment::ordered_pattern order, patterns repeated `frequency` times"""
        st = np.ascontiguousarray(site_states, dtype=np.uint8)
        assert st.ndim == 2 and st.shape[0] == self.ntaxa
        self._pars_sites = st
        self._ck(self.lib.lktree_set_parsimony_sites(self.h, st.shape[1], st.ctypes.data_as(c_uint8_p), int(aln_nsite), device))

    def computeParsimony(self):
        v = C.c_int32()
        self._ck(self.lib.lktree_compute_parsimony(self.h, C.byref(v)))
        return v.value

    def computeParsimonyBranch(self, node1, node2):
        """(parsimony score of the tree, substitutions on the branch) - PhyloTree::computeParsimonyBranch with branch_subst"""
        v, b = C.c_int32(), C.c_int32()
        self._ck(self.lib.lktree_compute_parsimony_branch(self.h, node1, node2, C.byref(v), C.byref(b)))
        return v.value, b.value

    def fixNegativeBranch(self, force=True):
        v = C.c_int32()
        self._ck(self.lib.lktree_fix_negative_branch(self.h, int(force), C.byref(v)))
        return v.value

    def computeParsimonyTree(self, taxon_order):
        """PhyloTree::computeParsimonyTree for the given addition order: replaces the tree, returns best_pars_score"""
        o = np.ascontiguousarray(taxon_order, dtype=np.int32)
        assert o.size == self.ntaxa
        v = C.c_int32()
        self._ck(self.lib.lktree_compute_parsimony_tree(self.h, o.ctypes.data_as(c_int32_p), C.byref(v)))
        return v.value

    def getBranches(self):
        cap = 4 * max(self.ntaxa, 2)
        pairs = np.zeros(2 * cap, dtype=np.int32)
        cnt = C.c_int32()
        self._ck(self.lib.lktree_get_branches(self.h, pairs.ctypes.data_as(c_int32_p), cap, C.byref(cnt)))
        return pairs[: 2 * cnt.value].reshape(-1, 2).copy()

    def getAdjacency(self):
        buf = C.create_string_buffer(256 * max(self.ntaxa, 4) + 4096)
        self._ck(self.lib.lktree_get_adjacency(self.h, buf, len(buf)))
        return buf.value.decode()

    def parsEngine(self):
        return self.lib.lktree_pars_engine(self.h)

    def parsLaunches(self):
        n, ms = C.c_uint64(), C.c_double()
        e = self.parsEngine()
        if not e:
            return 0, 0.0
        self.lib.lkpars_launch_count(e, C.byref(n), C.byref(ms))
        return n.value, ms.value

    def branchVector(self, frm, to):
        vec, comp = C.c_int32(), C.c_int32()
        self._ck(self.lib.lktree_branch_vector(self.h, frm, to, C.byref(vec), C.byref(comp)))
        return vec.value, bool(comp.value)

    def rootBranch(self):
        a, b = C.c_int32(), C.c_int32()
        self._ck(self.lib.lktree_root_branch(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def tipPartialLh(self):
        cap = 1 << 20
        out = np.zeros(cap, dtype=np.float64)
        cnt = C.c_int64()
        self._ck(self.lib.lktree_fetch_tip_partial_lh(self.h, _dp(out), cap, C.byref(cnt)))
        return out[: cnt.value].copy()

    def ptnInvar(self):
        out = np.zeros(self.nptn + self.nunobs + 1, dtype=np.float64)
        self._ck(self.lib.lktree_fetch_ptn_invar(self.h, _dp(out)))
        return out[: self.nptn + self.nunobs]

    def numPartialLhComputations(self):
        return int(self.lib.lktree_num_partial_lh_computations(self.h))

    # ---- engine-level access (shard i) --------------------------------------------------------------------
    def engine(self, i=0):
        e = self.lib.lktree_engine(self.h, i)
        if not e:
            raise LkError("no engine %d (initializeAllPartialLh first)" % i)
        return e

    def _eck(self, e, rc):
        if rc != 0:
            raise LkError(self.lib.lkcuda_last_error(e).decode() + " (status %d)" % rc)

    def numShards(self):
        n = 0
        while self.lib.lktree_engine(self.h, n):
            n += 1
        return n

    def shardPatterns(self, i):
        b, e = C.c_int64(), C.c_int64()
        self._ck(self.lib.lktree_shard_patterns(self.h, i, C.byref(b), C.byref(e)))
        return b.value, e.value

    def _fetch_shards(self, width, dtype, fetch):
        """concatenate the in-process shards: observed patterns in order, then the (replicated) unobserved ones"""
        obs, unobs = [], None
        for i in range(self.numShards()):
            b, en = self.shardPatterns(i)
            npat = (en - b) + self.nunobs
            buf = np.zeros(max(npat * width, 1), dtype=dtype)
            fetch(self.engine(i), buf)
            buf = buf[: npat * width].reshape(npat, width)
            obs.append(buf[: en - b])
            if unobs is None:
                unobs = buf[en - b:]
        return np.concatenate(obs + [unobs], axis=0)

    def fetchVector(self, vec):
        lh = self._fetch_shards(self.block, np.float64,
                                lambda e, buf: self._eck(e, self.lib.lkcuda_fetch_vector(e, vec, _dp(buf), None)))
        sc = self._fetch_shards(1, np.int16,
                                lambda e, buf: self._eck(e, self.lib.lkcuda_fetch_vector(e, vec, None, buf.ctypes.data_as(c_int16_p))))
        return lh, sc[:, 0].copy()

    def fetchTheta(self):
        return self._fetch_shards(self.block, np.float64, lambda e, buf: self._eck(e, self.lib.lkcuda_fetch_theta(e, _dp(buf))))

    def vectorScaleFactor(self, vec):
        tot = 0.0
        for i in range(self.numShards()):
            e = self.engine(i)
            v = C.c_double()
            self._eck(e, self.lib.lkcuda_vector_scale_factor(e, vec, C.byref(v)))
            tot += v.value
        return tot

    def setResultBuffer(self, dev_ptr, shard=0):
        """keep the 8-double result block in caller-owned device memory (e.g. a torch tensor's data_ptr()) for an in-place all-reduce"""
        e = self.engine(shard)
        self._eck(e, self.lib.lkcuda_set_result_buffer(e, C.c_void_p(dev_ptr)))

    def launchCount(self, shard=0):
        e = self.engine(shard)
        v = C.c_uint64()
        self._eck(e, self.lib.lkcuda_launch_count(e, C.byref(v)))
        return v.value

    def algorithmicBytes(self, shard=0):
        e = self.engine(shard)
        v = C.c_uint64()
        self._eck(e, self.lib.lkcuda_algorithmic_bytes(e, C.byref(v)))
        return v.value

    def memoryBytes(self, shard=0):
        e = self.engine(shard)
        v = C.c_uint64()
        self._eck(e, self.lib.lkcuda_memory_bytes(e, C.byref(v)))
        return v.value

    def resetCounters(self, shard=0):
        e = self.engine(shard)
        self._eck(e, self.lib.lkcuda_reset_counters(e))

    def setProfiling(self, on, shard=0):
        e = self.engine(shard)
        self._eck(e, self.lib.lkcuda_set_profiling(e, int(on)))

    def kernelStats(self, shard=0):
        """{kind: (launches, algorithmic bytes, device ms)}"""
        e = self.engine(shard)
        out = {}
        for k, name in enumerate(KIND_NAMES):
            n, b, ms = C.c_uint64(), C.c_uint64(), C.c_double()
            self._eck(e, self.lib.lkcuda_kernel_stats(e, k, C.byref(n), C.byref(b), C.byref(ms)))
            out[name] = (n.value, b.value, ms.value)
        return out

    def kernelFlops(self, shard=0):
        """{kind: algorithmic FP64 operations}"""
        e = self.engine(shard)
        out = {}
        for k, name in enumerate(KIND_NAMES):
            f = C.c_uint64()
            self._eck(e, self.lib.lkcuda_kernel_flops(e, k, C.byref(f)))
            out[name] = f.value
        return out

    def resultDevicePtr(self, shard=0):
        e = self.engine(shard)
        p = C.c_void_p()
        self._eck(e, self.lib.lkcuda_result_device_ptr(e, C.byref(p)))
        return p.value

    def synchronize(self, shard=0):
        e = self.engine(shard)
        self._eck(e, self.lib.lkcuda_synchronize(e))
