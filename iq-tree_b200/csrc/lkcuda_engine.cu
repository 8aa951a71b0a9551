/*
 * lkcuda_engine.cu — life cycle, inputs, host-visible mirrors and instrumentation of the sm_100a
 * likelihood engine (C ABI: include/lkcuda.h).  The kernels live in lkcuda_partial.cu and
 * lkcuda_reduce.cu.  There is no CPU fallback: without a CUDA device lkcuda_create fails with
 * LKCUDA_ERR_NODEVICE.
 *
 * Replaces the buffer management of PhyloTree::initializeAllPartialLh / deleteAllPartialLh
 * (phylotree.cpp:667-716, 718-760, 834-987): one arena for all partial-likelihood vectors, one for the
 * scale_num rows, plus _pattern_lh, _pattern_lh_cat, theta_all, ptn_freq, ptn_invar.
 */
#include "lkcuda_traverse.cuh"
#include <stdarg.h>
#include <stdlib.h>
#include <chrono>
#include <new>

static thread_local char g_create_err[512] = "";     /* engines are created concurrently from several host threads (phylosupertree.cpp:969-976) */

namespace lkc {

int fail(lkcuda_engine *e, int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(e ? e->err : g_create_err, 512, fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(lkcuda_engine *e, cudaError_t rc, const char *what) {
    return fail(e, rc == cudaErrorMemoryAllocation ? LKCUDA_ERR_NOMEM : LKCUDA_ERR_CUDA, "%s: %s", what, cudaGetErrorString(rc));
}

static cudaEvent_t get_event(lkcuda_engine *e) {
    cudaEvent_t ev = NULL;
    if (!e->event_pool.empty()) { ev = e->event_pool.back(); e->event_pool.pop_back(); return ev; }
    cudaEventCreate(&ev);
    return ev;
}

void launch_begin(lkcuda_engine *e, int kind) {
    if (!e->profiling) return;
    PendingEvent pe;
    pe.kind = kind; pe.start = get_event(e); pe.stop = get_event(e);
    cudaEventRecord(pe.start, e->stream);
    e->pending.push_back(pe);
}

int launch_end(lkcuda_engine *e, int kind, uint64_t algorithmic_bytes, const char *what) {
    cudaError_t rc = cudaGetLastError();
    if (e->profiling && !e->pending.empty()) cudaEventRecord(e->pending.back().stop, e->stream);
    e->launches++; e->bytes += algorithmic_bytes;
    e->kind_launches[kind]++; e->kind_bytes[kind] += algorithmic_bytes;
    if (rc != cudaSuccess) return cuda_fail(e, rc, what);
    if (e->pending.size() > 4096) drain_events(e);
    return LKCUDA_OK;
}

/* fold finished event pairs into kind_ms (synchronises on them) */
void drain_events(lkcuda_engine *e) {
    for (size_t i = 0; i < e->pending.size(); i++) {
        PendingEvent &pe = e->pending[i];
        float ms = 0.f;
        if (cudaEventSynchronize(pe.stop) == cudaSuccess && cudaEventElapsedTime(&ms, pe.start, pe.stop) == cudaSuccess)
            e->kind_ms[pe.kind] += ms;
        e->event_pool.push_back(pe.start); e->event_pool.push_back(pe.stop);
    }
    e->pending.clear();
}

int states_barrier(lkcuda_engine *e) {
    if (!e->upload_pending) return LKCUDA_OK;
    LK_CUDA(e, cudaStreamWaitEvent(e->stream, e->ev_upload_done, 0));
    e->upload_pending = false;
    return LKCUDA_OK;
}

int upload_verdict(lkcuda_engine *e) {
    if (!e->upload_check) return LKCUDA_OK;
    LK_CUDA(e, cudaEventSynchronize(e->ev_upload_done));
    e->upload_check = false;
    if (e->h_upload_bad[0] != 0) { e->states_set = false; return fail(e, LKCUDA_ERR_ARG, "tip state out of range"); }
    return LKCUDA_OK;
}

}  // namespace lkc

using namespace lkc;

static long long round_up(long long x, long long m) { return (x + m - 1) / m * m; }

template <typename T>
static cudaError_t dev_alloc(lkcuda_engine *e, T **p, size_t count) {
    size_t bytes = (count ? count : 1) * sizeof(T);
    cudaError_t rc = cudaMalloc((void **)p, bytes);
    if (rc == cudaSuccess) e->total_dev_bytes += bytes;
    return rc;
}

/* flags any state code above state_unknown: gridDim.y = taxa (rows of the matrix), 16 bytes per thread and step, byte-wise SIMD compare */
__global__ void validate_states_kernel(const uint8_t *states, int ntaxa, long long stride, long long np, int su, unsigned int *flag) {
    const unsigned int su4 = (unsigned int)su * 0x01010101u;
    bool bad = false;
    for (int t = blockIdx.y; t < ntaxa; t += gridDim.y) {
    const uint8_t *row = states + (long long)t * stride;
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 16; i < np; i += (long long)gridDim.x * blockDim.x * 16) {
        const uint4 w = *reinterpret_cast<const uint4 *>(row + i);
        if (i + 16 <= np) bad |= (__vcmpgtu4(w.x, su4) | __vcmpgtu4(w.y, su4) | __vcmpgtu4(w.z, su4) | __vcmpgtu4(w.w, su4)) != 0u;
        else {
            const unsigned int x[4] = {w.x, w.y, w.z, w.w};
            for (int k = 0; k < (int)(np - i); k++) bad |= (int)((x[k >> 2] >> ((k & 3) * 8)) & 0xff) > su;
        }
    }
    }
    if (bad) atomicOr(flag, 1u);
}

static void launch_validate(lkcuda_engine *e, cudaStream_t s, unsigned int *flag) {
    const DevView &v = e->v;
    long long per_row = (v.NP / 16 + 255) / 256;
    int gx = (int)(per_row < 1 ? 1 : (per_row > 64 ? 64 : per_row));
    validate_states_kernel<<<dim3(gx, e->cfg.ntaxa < 65535 ? e->cfg.ntaxa : 65535), 256, 0, s>>>(e->d_states, e->cfg.ntaxa, v.st_stride, v.NP, e->cfg.state_unknown, flag);
}

extern "C" {

int lkcuda_device_count(int32_t *count) {
    if (!count) return LKCUDA_ERR_ARG;
    int n = 0;
    cudaError_t rc = cudaGetDeviceCount(&n);
    if (rc != cudaSuccess) { *count = 0; cudaGetLastError(); return LKCUDA_ERR_NODEVICE; }
    *count = n;
    return LKCUDA_OK;
}

const char *lkcuda_last_error(const lkcuda_engine *e) { return e ? e->err : g_create_err; }

int lkcuda_create(const lkcuda_config *cfg, lkcuda_engine **out) {
    if (!cfg || !out) return fail(NULL, LKCUDA_ERR_ARG, "null argument");
    *out = NULL;
    if ((cfg->flags & LKCUDA_FLAG_SITEMODEL) && (cfg->nstates > 32 || cfg->nunobs > 0))
        return fail(NULL, LKCUDA_ERR_ARG, "site-specific models: at most 32 states, no +ASC");
    if (cfg->nstates < 1 || cfg->nblocks < 1 || cfg->neigen < 1 || cfg->ntaxa < 1 || cfg->nptn < 0 ||
        cfg->nunobs < 0 || cfg->nvec < 0 || cfg->state_unknown < cfg->nstates || cfg->state_unknown > 255)
        return fail(NULL, LKCUDA_ERR_ARG, "invalid configuration");
    if ((long long)cfg->nstates * cfg->nblocks > (1 << 20)) return fail(NULL, LKCUDA_ERR_ARG, "block too large");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        return fail(NULL, LKCUDA_ERR_NODEVICE, "no CUDA device available (this engine has no CPU path)");
    }
    if (cfg->device < 0 || cfg->device >= ndev) return fail(NULL, LKCUDA_ERR_ARG, "device ordinal %d out of range (%d devices)", cfg->device, ndev);
    cudaError_t rc = cudaSetDevice(cfg->device);
    if (rc != cudaSuccess) return cuda_fail(NULL, rc, "cudaSetDevice");

    lkcuda_engine *e = new (std::nothrow) lkcuda_engine();
    if (!e) return fail(NULL, LKCUDA_ERR_NOMEM, "out of host memory");
    /* lkcuda_engine holds std::vector members: value-initialise the POD part by hand */
    e->aux = NULL;
    e->cfg = *cfg; e->device = cfg->device; e->stream = NULL; e->own_stream = false;
    e->d_states = NULL; e->d_freq = e->d_invar = NULL;
    e->d_eval = e->d_evec = e->d_inv = e->d_rate = e->d_weight = e->d_tip = NULL; e->d_eigen = NULL;
    e->d_arena = NULL; e->d_scale = NULL; e->d_theta = NULL; e->d_theta_sc = NULL; e->d_theta_meta = NULL;
    e->dstream = NULL; e->ev_plh_ready = NULL; e->ev_plh_done = NULL; e->plh_pending = false;
    e->d_plh = e->d_plhcat = e->d_plh_out = NULL; e->d_red = NULL; e->d_counter = NULL; e->d_result = NULL; e->h_result = NULL; e->d_lens = NULL; e->h_lens = NULL;
    e->d_newton_out = NULL; e->h_newton_out = NULL; e->d_newton_ctl = NULL; e->newton_max_grid = 0;
    e->d_ptn_model = NULL; e->d_state_mask = NULL;
    e->h_mapped = NULL; e->result_seq = 0; e->d_mbox = NULL; e->d_peer_tab = NULL; e->peer_seq = 0;
    memset(e->peer_opened, 0, sizeof(e->peer_opened));
    e->d_tables = NULL; e->table_slots_cap = 0;
    for (int i = 0; i < lkcuda_engine::RING; i++) {
        e->h_ops[i] = e->d_ops[i] = NULL; e->h_child[i] = e->d_child[i] = NULL;
        e->ring_cap_ops[i] = e->ring_cap_child[i] = 0; e->ring_ev[i] = NULL; e->ring_used[i] = false;
    }
    e->ring_next = 0; e->red_blocks = 0; e->trav_smem_set = 48 * 1024; e->fuse_state = 0;
    e->d_inv_frags = NULL; e->mma_smem_set = 48 * 1024;
    { const char *r = getenv("LKCUDA_NO_MMA"); e->trav_disable_mma = r && atoi(r) != 0; }
    { const char *r = getenv("LKCUDA_TRV_MAXSLOTS"); e->trav_maxslots = r ? atoi(r) : -1; }
    e->theta_ready = false; e->theta_dad = e->theta_node = 0;
    e->cstream = NULL; e->ev_upload_done = e->ev_main_mark = NULL; e->d_upload_flag = NULL; e->h_upload_vals = e->h_upload_bad = NULL;
    e->upload_base = 0; e->upload_chunks = 0; e->upload_chunk_pat = 0; e->upload_pending = e->upload_check = false;
    e->model_set = e->states_set = e->weights_set = false;
    e->launches = e->bytes = 0; e->profiling = false; e->total_dev_bytes = 0;
    memset(e->kind_ms, 0, sizeof(e->kind_ms)); memset(e->kind_launches, 0, sizeof(e->kind_launches)); memset(e->kind_bytes, 0, sizeof(e->kind_bytes)); memset(e->kind_flops, 0, sizeof(e->kind_flops));
    e->err[0] = 0;

    DevView &v = e->v;
    memset(&v, 0, sizeof(v));
    v.S = cfg->nstates; v.K = cfg->nblocks; v.E = cfg->neigen; v.B = v.S * v.K; v.SU = cfg->state_unknown;
    v.flags = cfg->flags; v.P = cfg->nptn; v.U = cfg->nunobs; v.NP = v.P + v.U; v.nsite = (double)cfg->nsite;
    v.vec_stride = round_up(v.NP * v.B, 32);     /* 256-byte multiple */
    v.sc_stride = round_up(v.NP, 64);            /* 128-byte multiple */
    v.st_stride = round_up(v.NP, 128);
    v.table_slot = round_up((long long)v.K * (v.S * v.S + 2) + (long long)(v.SU + 1) * v.B + v.SU + 4, 4);
    if (v.S == 20 && v.K <= 4) {    /* protein MMA kernel: K fragment-ordered matrices (3*3*2*32 doubles each) + tip table with 24-double rows */
        long long need = (long long)v.K * 576 + (long long)(v.SU + 1) * v.K * 24;
        if (need > v.table_slot) v.table_slot = round_up(need, 4);
    }   /* + 2: padded matrices of the DNA traversal kernel */
    { long long need = mmak_table_slot(v); if (need > v.table_slot) v.table_slot = round_up(need, 4); }   /* codon / wide-block protein MMA kernel */

    cudaDeviceProp prop;
    rc = cudaGetDeviceProperties(&prop, cfg->device);
    e->num_sms = rc == cudaSuccess ? prop.multiProcessorCount : 148;
    e->red_blocks = e->num_sms * 16;

    bool ok = true;
#define ALLOC(ptr, count) do { if (ok) { rc = dev_alloc(e, &(ptr), (size_t)(count)); if (rc != cudaSuccess) ok = false; } } while (0)
    if (ok) { rc = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking); ok = rc == cudaSuccess; e->own_stream = ok; }
    size_t nvec = (size_t)(cfg->nvec ? cfg->nvec : 1);
    ALLOC(e->d_states, (size_t)cfg->ntaxa * v.st_stride + 4096);   /* + slack: the traversal kernel bulk-copies whole tiles of a row */
    ALLOC(e->d_freq, v.P + 4); ALLOC(e->d_invar, v.NP + 4);
    ALLOC(e->d_eval, v.E * v.S); ALLOC(e->d_evec, (size_t)v.E * v.S * v.S); ALLOC(e->d_inv, (size_t)v.E * v.S * v.S);
    ALLOC(e->d_rate, v.K); ALLOC(e->d_weight, v.K); ALLOC(e->d_eigen, v.K);
    ALLOC(e->d_tip, (cfg->flags & LKCUDA_FLAG_SITEMODEL) ? 4 : (size_t)(v.SU + 1) * v.E * v.S);
    ALLOC(e->d_arena, nvec * (size_t)v.vec_stride);
    ALLOC(e->d_scale, nvec * (size_t)v.sc_stride);
    ALLOC(e->d_theta, (size_t)v.vec_stride);
    ALLOC(e->d_theta_sc, v.U + 4); ALLOC(e->d_theta_meta, 8);
    ALLOC(e->d_plh, v.NP + 4); ALLOC(e->d_plhcat, (size_t)v.P * v.K + 4);
    ALLOC(e->d_red, (size_t)e->red_blocks * RED_SLOTS); ALLOC(e->d_counter, 4); ALLOC(e->d_result, LKCUDA_NRESULT);
    ALLOC(e->d_newton_out, 8);
    ALLOC(e->d_lens, 5 * LKCUDA_NLENSLOT);
    ALLOC(e->d_upload_flag, 4);
#undef ALLOC
    if (ok) { rc = cudaMallocHost((void **)&e->h_result, sizeof(double) * LKCUDA_NRESULT); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaMallocHost((void **)&e->h_newton_out, sizeof(double) * 8); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaMallocHost((void **)&e->h_lens, sizeof(double) * 5 * LKCUDA_NLENSLOT); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaMemsetAsync(e->d_lens, 0, sizeof(double) * 5 * LKCUDA_NLENSLOT, e->stream); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaHostAlloc((void **)&e->h_mapped, sizeof(double) * 2 * LKCUDA_NRESULT, cudaHostAllocMapped); ok = rc == cudaSuccess; }
    if (ok) memset(e->h_mapped, 0, sizeof(double) * 2 * LKCUDA_NRESULT);
    if (ok) { rc = cudaMallocHost((void **)&e->h_upload_vals, sizeof(unsigned int) * (lkcuda_engine::UPLOAD_MAX_CHUNKS + 1)); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaMallocHost((void **)&e->h_upload_bad, sizeof(unsigned int) * 4); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaStreamCreateWithFlags(&e->cstream, cudaStreamNonBlocking); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaEventCreateWithFlags(&e->ev_upload_done, cudaEventDisableTiming); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaEventCreateWithFlags(&e->ev_main_mark, cudaEventDisableTiming); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaMemsetAsync(e->d_upload_flag, 0, 4 * sizeof(unsigned int), e->stream); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaMemsetAsync(e->d_counter, 0, 4 * sizeof(unsigned int), e->stream); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaMemsetAsync(e->d_red, 0, (size_t)e->red_blocks * RED_SLOTS * sizeof(uint4), e->stream); ok = rc == cudaSuccess; }    /* tags of generation 0 */
    if (ok) { rc = cudaMemsetAsync(e->d_result, 0, sizeof(double) * LKCUDA_NRESULT, e->stream); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaMemsetAsync(e->d_theta_meta, 0, sizeof(double) * 8, e->stream); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaMemsetAsync(e->d_states, 0, (size_t)cfg->ntaxa * v.st_stride + 4096, e->stream); ok = rc == cudaSuccess; }
    {   /* test knob: fill everything a kernel must write before anything reads it with NaN bit patterns, so that a pattern no kernel
           covered cannot pass for computed because the allocation happened to hold an earlier run's values */
        static const bool poison = getenv("LKCUDA_POISON") && atoi(getenv("LKCUDA_POISON")) != 0;
        if (poison && ok) {
            cudaMemsetAsync(e->d_arena, 0xff, nvec * (size_t)v.vec_stride * 8, e->stream);
            cudaMemsetAsync(e->d_scale, 0x7f, nvec * (size_t)v.sc_stride * 2, e->stream);
            cudaMemsetAsync(e->d_theta, 0xff, (size_t)v.vec_stride * 8, e->stream);
            cudaMemsetAsync(e->d_plh, 0xff, ((size_t)v.NP + 4) * 8, e->stream);
        }
    }
    if (ok) { rc = cudaMemsetAsync(e->d_scale, 0, nvec * (size_t)v.sc_stride * sizeof(int16_t), e->stream); ok = rc == cudaSuccess; }
    if (ok) { rc = cudaStreamSynchronize(e->stream); ok = rc == cudaSuccess; }
    if (!ok) {
        int code = cuda_fail(NULL, rc, "engine allocation");
        lkcuda_destroy(e);
        return code;
    }
    v.states = e->d_states; v.freq = e->d_freq; v.invar = e->d_invar;
    v.eval = e->d_eval; v.evec = e->d_evec; v.inv_evec = e->d_inv; v.blk_rate = e->d_rate; v.blk_weight = e->d_weight;
    v.tip = e->d_tip; v.blk_eigen = e->d_eigen;
    e->own_vec_stride = v.vec_stride; e->own_sc_stride = v.sc_stride; e->own_nvec = cfg->nvec;
    v.arena = e->d_arena; v.scale = e->d_scale; v.theta = e->d_theta; v.theta_sc = e->d_theta_sc; v.theta_meta = e->d_theta_meta;
    v.pattern_lh = e->d_plh; v.pattern_lh_cat = e->d_plhcat;
    v.red = e->d_red; v.counter = e->d_counter; v.result = e->d_result; v.lens = e->d_lens;
    v.tables = NULL;
    {   /* device-side address of the host-mapped result block (identical under UVA; asked for all the same) */
        void *dp = NULL;
        static const bool no_mapped = getenv("LKCUDA_NO_MAPPED_RESULT") && atoi(getenv("LKCUDA_NO_MAPPED_RESULT")) != 0;     /* measurement knob: results by copy */
        if (!no_mapped && cudaHostGetDevicePointer(&dp, e->h_mapped, 0) == cudaSuccess) v.host_result = (double *)dp;
        else { cudaGetLastError(); v.host_result = NULL; }
    }
    v.seq = 0; v.peer_world = 1; v.peer_rank = 0; v.peer_seq = 0; v.peer_mbox = NULL;
    *out = e;
    return LKCUDA_OK;
}

int lkcuda_destroy(lkcuda_engine *e) {
    if (!e) return LKCUDA_OK;
    cudaSetDevice(e->device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    if (e->cstream) { cudaStreamSynchronize(e->cstream); cudaStreamDestroy(e->cstream); }
    if (e->ev_upload_done) cudaEventDestroy(e->ev_upload_done);
    if (e->ev_main_mark) cudaEventDestroy(e->ev_main_mark);
    cudaFree(e->d_upload_flag);
    if (e->h_upload_vals) cudaFreeHost(e->h_upload_vals);
    if (e->h_upload_bad) cudaFreeHost(e->h_upload_bad);
    drain_events(e);
    aux_destroy(e);
    for (size_t i = 0; i < e->event_pool.size(); i++) cudaEventDestroy(e->event_pool[i]);
    cudaFree(e->d_states); cudaFree(e->d_freq); cudaFree(e->d_invar);
    cudaFree(e->d_eval); cudaFree(e->d_evec); cudaFree(e->d_inv); cudaFree(e->d_rate); cudaFree(e->d_weight); cudaFree(e->d_eigen);
    cudaFree(e->d_tip); cudaFree(e->d_arena); cudaFree(e->d_scale); cudaFree(e->d_theta); cudaFree(e->d_theta_sc); cudaFree(e->d_theta_meta);
    if (e->dstream) { cudaStreamSynchronize(e->dstream); cudaStreamDestroy(e->dstream); cudaEventDestroy(e->ev_plh_ready); cudaEventDestroy(e->ev_plh_done); }
    cudaFree(e->d_plh); cudaFree(e->d_plhcat); cudaFree(e->d_plh_out); cudaFree(e->d_red); cudaFree(e->d_counter); cudaFree(e->d_result); cudaFree(e->d_lens); if (e->h_lens) cudaFreeHost(e->h_lens);
    cudaFree(e->d_newton_out); cudaFree(e->d_newton_ctl); cudaFree(e->d_tables); cudaFree(e->d_inv_frags);
    if (e->h_result) cudaFreeHost(e->h_result);
    if (e->h_newton_out) cudaFreeHost(e->h_newton_out);
    if (e->h_mapped) cudaFreeHost(e->h_mapped);
    for (int q = 0; q < LKCUDA_MAX_PEERS; q++) if (e->peer_opened[q]) cudaIpcCloseMemHandle(e->peer_opened[q]);
    cudaFree(e->d_mbox); cudaFree(e->d_peer_tab); cudaFree(e->d_ptn_model); cudaFree(e->d_state_mask);
    for (int i = 0; i < lkcuda_engine::RING; i++) {
        if (e->h_ops[i]) cudaFreeHost(e->h_ops[i]);
        if (e->h_child[i]) cudaFreeHost(e->h_child[i]);
        cudaFree(e->d_ops[i]); cudaFree(e->d_child[i]);
        if (e->ring_ev[i]) cudaEventDestroy(e->ring_ev[i]);
    }
    if (e->own_stream && e->stream) cudaStreamDestroy(e->stream);
    cudaGetLastError();
    delete e;
    return LKCUDA_OK;
}

/* ---- externally owned arenas (PhyloSuperTreePlen::initializeAllPartialLh, phylosupertreeplen.cpp:1742-1871, hands every
   partition a window of ONE shared arena; ModelMixture::optimizeWithEM aliases a tree's arena into a temporary tree,
   model/modelmixture.cpp:1427-1429, 1524-1526) ---- */
int lkcuda_device_malloc(int32_t device, uint64_t bytes, void **ptr) {
    if (!ptr) return LKCUDA_ERR_ARG;
    *ptr = NULL;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) { cudaGetLastError(); return fail(NULL, LKCUDA_ERR_NODEVICE, "no CUDA device available"); }
    if (device < 0 || device >= ndev) return fail(NULL, LKCUDA_ERR_ARG, "device ordinal %d out of range", device);
    cudaError_t rc = cudaSetDevice(device);
    if (rc == cudaSuccess) rc = cudaMalloc(ptr, bytes ? (size_t)bytes : 1);
    if (rc == cudaSuccess) rc = cudaMemset(*ptr, 0, bytes ? (size_t)bytes : 1);
    if (rc != cudaSuccess) return cuda_fail(NULL, rc, "lkcuda_device_malloc");
    return LKCUDA_OK;
}

int lkcuda_device_free(int32_t device, void *ptr) {
    if (!ptr) return LKCUDA_OK;
    cudaError_t rc = cudaSetDevice(device);
    if (rc == cudaSuccess) rc = cudaDeviceSynchronize();
    if (rc == cudaSuccess) rc = cudaFree(ptr);
    if (rc != cudaSuccess) return cuda_fail(NULL, rc, "lkcuda_device_free");
    return LKCUDA_OK;
}

int lkcuda_host_register(lkcuda_engine *e, void *ptr, uint64_t bytes) {
    LK_ENTER(e);
    if (!ptr || !bytes) return fail(e, LKCUDA_ERR_ARG, "lkcuda_host_register: null array");
    cudaError_t rc = cudaHostRegister(ptr, (size_t)bytes, cudaHostRegisterDefault);
    if (rc == cudaErrorHostMemoryAlreadyRegistered) { cudaGetLastError(); return LKCUDA_OK; }
    if (rc != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, rc, "cudaHostRegister"); }
    return LKCUDA_OK;
}

int lkcuda_host_unregister(lkcuda_engine *e, void *ptr) {
    LK_ENTER(e);
    if (!ptr) return LKCUDA_OK;
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    cudaError_t rc = cudaHostUnregister(ptr);
    if (rc == cudaErrorHostMemoryNotRegistered) { cudaGetLastError(); return LKCUDA_OK; }
    if (rc != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, rc, "cudaHostUnregister"); }
    return LKCUDA_OK;
}

int lkcuda_arena_layout(const lkcuda_engine *e, int64_t *vec_stride, int64_t *scale_stride) {
    if (!e) return LKCUDA_ERR_ARG;
    if (vec_stride) *vec_stride = e->v.vec_stride;
    if (scale_stride) *scale_stride = e->v.sc_stride;
    return LKCUDA_OK;
}

int lkcuda_attach_arena(lkcuda_engine *e, void *partial_lh, int64_t vec_stride, void *scale_num, int64_t scale_stride, int32_t nvec) {
    LK_ENTER(e);
    DevView &v = e->v;
    if (!partial_lh && !scale_num) {            /* detach: back to the engine's own arena (sized by cfg.nvec at creation) */
        LK_CUDA(e, cudaStreamSynchronize(e->stream));
        v.arena = e->d_arena; v.scale = e->d_scale;
        v.vec_stride = e->own_vec_stride; v.sc_stride = e->own_sc_stride; e->cfg.nvec = e->own_nvec;
        e->theta_ready = false;
        return LKCUDA_OK;
    }
    if (!partial_lh || !scale_num || nvec < 0) return fail(e, LKCUDA_ERR_ARG, "lkcuda_attach_arena: null base address");
    if (vec_stride < v.NP * v.B || scale_stride < v.NP) return fail(e, LKCUDA_ERR_ARG, "lkcuda_attach_arena: stride shorter than one vector");
    /* the kernels move whole 32-byte sectors: vectors must start on 32-byte boundaries */
    if (((uintptr_t)partial_lh & 31) || (v.B % 4 == 0 && vec_stride % 4 != 0) || ((uintptr_t)scale_num & 1))
        return fail(e, LKCUDA_ERR_ARG, "lkcuda_attach_arena: base must be 32-byte aligned and the stride a multiple of 4 doubles");
    const void *both[2] = {partial_lh, scale_num};
    for (int k = 0; k < 2; k++) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, both[k]) != cudaSuccess || at.type != cudaMemoryTypeDevice || at.device != e->device) {
            cudaGetLastError();
            return fail(e, LKCUDA_ERR_ARG, "lkcuda_attach_arena: %s is not device memory of device %d", k ? "scale_num" : "partial_lh", e->device);
        }
    }
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    v.arena = (double *)partial_lh; v.scale = (int16_t *)scale_num;
    v.vec_stride = vec_stride; v.sc_stride = scale_stride; e->cfg.nvec = nvec;
    e->theta_ready = false;
    return LKCUDA_OK;
}

int lkcuda_memory_bytes(const lkcuda_engine *e, uint64_t *bytes) {
    if (!e || !bytes) return LKCUDA_ERR_ARG;
    *bytes = e->total_dev_bytes;
    return LKCUDA_OK;
}

int lkcuda_set_stream(lkcuda_engine *e, void *s) {
    LK_ENTER(e);
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    drain_events(e);
    if (e->own_stream) { cudaStreamDestroy(e->stream); e->own_stream = false; }
    if (s) e->stream = (cudaStream_t)s;
    else { LK_CUDA(e, cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking)); e->own_stream = true; }
    return LKCUDA_OK;
}

int lkcuda_synchronize(lkcuda_engine *e) {
    LK_ENTER(e);
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    return upload_verdict(e);
}

int lkcuda_set_tip_states_async(lkcuda_engine *e, const uint8_t *states, int32_t nchunks) {
    LK_ENTER(e);
    if (!states) return fail(e, LKCUDA_ERR_ARG, "null tip states");
    const DevView &v = e->v;
    if (nchunks < 1) nchunks = 1;
    if (nchunks > lkcuda_engine::UPLOAD_MAX_CHUNKS) nchunks = lkcuda_engine::UPLOAD_MAX_CHUNKS;
    /* the previous upload's flag sources and verdict are reused: it must be over (it is, unless uploads are issued back to back) */
    if (e->upload_check || e->upload_pending) LK_CUDA(e, cudaEventSynchronize(e->ev_upload_done));
    if (e->upload_check) { int rc = upload_verdict(e); if (rc) return rc; }
    /* chunk boundaries on 128-pattern multiples: whole 16-byte granules for the kernel's bulk copies */
    long long chunk = ((v.NP + nchunks - 1) / nchunks + 127) / 128 * 128;
    if (chunk < 128) chunk = 128;
    nchunks = v.NP > 0 ? (int)((v.NP + chunk - 1) / chunk) : 0;
    /* kernels already queued on the engine stream still read the old matrix */
    LK_CUDA(e, cudaEventRecord(e->ev_main_mark, e->stream));
    LK_CUDA(e, cudaStreamWaitEvent(e->cstream, e->ev_main_mark, 0));
    e->upload_base += (unsigned int)e->upload_chunks;       /* the counter only ever grows: no reset to race with */
    e->upload_chunks = nchunks; e->upload_chunk_pat = chunk;
    for (int c = 0; c < nchunks; c++) {
        const long long b = (long long)c * chunk, w = (b + chunk <= v.NP ? chunk : v.NP - b);
        LK_CUDA(e, cudaMemcpy2DAsync(e->d_states + b, (size_t)v.st_stride, states + b, (size_t)v.NP, (size_t)w, (size_t)e->cfg.ntaxa,
                                     cudaMemcpyHostToDevice, e->cstream));
        e->h_upload_vals[c] = e->upload_base + (unsigned int)(c + 1);
        LK_CUDA(e, cudaMemcpyAsync(e->d_upload_flag, &e->h_upload_vals[c], sizeof(unsigned int), cudaMemcpyHostToDevice, e->cstream));
    }
    if (v.NP > 0) {
        LK_CUDA(e, cudaMemsetAsync(e->d_upload_flag + 1, 0, sizeof(unsigned int), e->cstream));
        launch_validate(e, e->cstream, e->d_upload_flag + 1);
        e->launches++; e->kind_launches[KIND_MISC]++;
        LK_CUDA(e, cudaGetLastError());
        LK_CUDA(e, cudaMemcpyAsync(e->h_upload_bad, e->d_upload_flag + 1, sizeof(unsigned int), cudaMemcpyDeviceToHost, e->cstream));
    } else e->h_upload_bad[0] = 0;
    LK_CUDA(e, cudaEventRecord(e->ev_upload_done, e->cstream));
    e->upload_pending = true; e->upload_check = v.NP > 0;
    e->states_set = true; e->theta_ready = false;
    return LKCUDA_OK;
}

int lkcuda_set_tip_states(lkcuda_engine *e, const uint8_t *states) {
    LK_ENTER(e);
    if (!states) return fail(e, LKCUDA_ERR_ARG, "null tip states");
    const DevView &v = e->v;
    if (e->upload_pending || e->upload_check) {      /* a pipelined upload is still in flight: finish it first */
        LK_CUDA(e, cudaEventSynchronize(e->ev_upload_done));
        e->upload_pending = false; e->upload_check = false;
    }
    if (v.NP > 0) {
        /* `states` may be pinned (then this is one DMA at link speed) or pageable caller memory */
        LK_CUDA(e, cudaMemcpy2DAsync(e->d_states, (size_t)v.st_stride, states, (size_t)v.NP, (size_t)v.NP, (size_t)e->cfg.ntaxa,
                                     cudaMemcpyHostToDevice, e->stream));
        /* range check on the device: the matrix is 100 MB at the BASELINE shape, a host scan would cost more than the copy */
        LK_CUDA(e, cudaMemsetAsync(e->d_counter + 2, 0, sizeof(unsigned int), e->stream));
        launch_begin(e, KIND_MISC);
        launch_validate(e, e->stream, e->d_counter + 2);
        int rc = launch_end(e, KIND_MISC, 0, "validate_states_kernel");
        if (rc) return rc;
        LK_CUDA(e, cudaMemcpyAsync(e->h_newton_out, e->d_counter + 2, sizeof(unsigned int), cudaMemcpyDeviceToHost, e->stream));
    }
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    if (v.NP > 0 && *(unsigned int *)e->h_newton_out != 0) { e->states_set = false; return fail(e, LKCUDA_ERR_ARG, "tip state out of range"); }
    e->states_set = true; e->theta_ready = false;
    return LKCUDA_OK;
}

int lkcuda_set_pattern_weights(lkcuda_engine *e, const double *ptn_freq, const double *ptn_invar) {
    LK_ENTER(e);
    if (!ptn_freq || !ptn_invar) return fail(e, LKCUDA_ERR_ARG, "null pattern weights");
    const DevView &v = e->v;
    if (v.P) LK_CUDA(e, cudaMemcpyAsync(e->d_freq, ptn_freq, (size_t)v.P * 8, cudaMemcpyHostToDevice, e->stream));
    if (v.NP) LK_CUDA(e, cudaMemcpyAsync(e->d_invar, ptn_invar, (size_t)v.NP * 8, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    e->weights_set = true; e->theta_ready = false;
    return LKCUDA_OK;
}

int lkcuda_set_model(lkcuda_engine *e, const double *eval, const double *evec, const double *inv_evec,
                     const int32_t *blk_eigen, const double *blk_rate, const double *blk_weight,
                     const double *tip_partial_lh) {
    LK_ENTER(e);
    const DevView &v = e->v;
    const bool sm = (v.flags & LKCUDA_FLAG_SITEMODEL) != 0;     /* blk_eigen and the tip table are not used: one eigensystem per pattern */
    if (!eval || !evec || !inv_evec || !blk_rate || !blk_weight || (!sm && (!blk_eigen || !tip_partial_lh)))
        return fail(e, LKCUDA_ERR_ARG, "null model argument");
    for (int k = 0; k < v.K && !sm; k++)
        if (blk_eigen[k] < 0 || blk_eigen[k] >= v.E) return fail(e, LKCUDA_ERR_ARG, "blk_eigen out of range");
    size_t S = (size_t)v.S, E = (size_t)v.E, K = (size_t)v.K;
    LK_CUDA(e, cudaMemcpyAsync(e->d_eval, eval, E * S * 8, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_evec, evec, E * S * S * 8, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_inv, inv_evec, E * S * S * 8, cudaMemcpyHostToDevice, e->stream));
    if (!sm) LK_CUDA(e, cudaMemcpyAsync(e->d_eigen, blk_eigen, K * 4, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_rate, blk_rate, K * 8, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaMemcpyAsync(e->d_weight, blk_weight, K * 8, cudaMemcpyHostToDevice, e->stream));
    if (!sm) LK_CUDA(e, cudaMemcpyAsync(e->d_tip, tip_partial_lh, (size_t)(v.SU + 1) * E * S * 8, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));   /* the sources are pageable caller memory */
    if (!sm && v.S == 4 && v.K == 4) {      /* parameter block of the B = 16 reductions (same products as build_val forms on the device) */
        for (int k = 0; k < 4; k++) {
            e->b16tab.w[k] = blk_weight[k]; e->b16tab.eig[k] = blk_eigen[k];
            for (int i = 0; i < 4; i++) e->b16tab.cof[4 * k + i] = eval[blk_eigen[k] * 4 + i] * blk_rate[k];
        }
    }
    e->model_set = true; e->theta_ready = false;
    return LKCUDA_OK;
}

int lkcuda_read_result(lkcuda_engine *e, double *result) {
    LK_ENTER(e);
    if (!result) return fail(e, LKCUDA_ERR_ARG, "null result");
    LK_CUDA(e, cudaMemcpyAsync(e->h_result, e->v.result, sizeof(double) * LKCUDA_NRESULT, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    memcpy(result, e->h_result, sizeof(double) * LKCUDA_NRESULT);
    return upload_verdict(e);
}

int lkcuda_wait_result(lkcuda_engine *e, double *result) {
    LK_ENTER(e);
    if (!result) return fail(e, LKCUDA_ERR_ARG, "null result");
    if (!e->v.host_result || e->result_seq == 0) return lkcuda_read_result(e, result);
    /* slot i = two 8-byte words {lo32 | tag << 32, hi32 | tag << 32}; tag = sequence number | error << 31 (post_result): a value is complete
       when both of its tags show the number, the block when all eight values are */
    volatile unsigned long long *slots = reinterpret_cast<volatile unsigned long long *>(e->h_mapped);
    const unsigned int want = e->result_seq;
    unsigned long long words[2 * LKCUDA_NRESULT];
    auto all_there = [&](unsigned long long &last) {
        for (int i = 2 * LKCUDA_NRESULT - 1; i >= 0; i--) {
            words[i] = slots[i]; last = words[i] >> 32;
            if (((unsigned int)last & 0x7fffffffu) != want) return false;
        }
        return true;
    };
    unsigned long long f = 0;
    if (!all_there(f)) {
        auto next_query = std::chrono::steady_clock::now() + std::chrono::microseconds(200);
        for (unsigned int spin = 0;; spin++) {
            if (all_there(f)) break;
#if defined(__x86_64__) || defined(__i386__)
            __builtin_ia32_pause();
#endif
            if ((spin & 63u) == 63u && std::chrono::steady_clock::now() >= next_query) {
                /* the stream drained (or died) without the number showing up: take the slow path, which also reports CUDA errors */
                cudaError_t q = cudaStreamQuery(e->stream);
                if (q != cudaErrorNotReady) {
                    if (q == cudaSuccess && all_there(f)) break;
                    if (q != cudaSuccess) return cuda_fail(e, q, "reduction kernel");
                    return lkcuda_read_result(e, result);
                }
                next_query = std::chrono::steady_clock::now() + std::chrono::microseconds(200);
            }
        }
    }
    for (int i = 0; i < LKCUDA_NRESULT; i++) {
        const unsigned long long bits = (words[2 * i] & 0xffffffffull) | (words[2 * i + 1] << 32);
        memcpy(&result[i], &bits, 8);
    }
    if ((f >> 31) & 1ull) return fail(e, LKCUDA_ERR_STATE, "cross-GPU all-reduce timed out: a peer never enqueued its reduction");
    return upload_verdict(e);
}

int lkcuda_peer_handle(lkcuda_engine *e, uint8_t *handle, void **local_ptr) {
    LK_ENTER(e);
    if (!e->d_mbox) {
        LK_CUDA(e, cudaMalloc((void **)&e->d_mbox, sizeof(PeerMbox)));     /* plain cudaMalloc: exportable through CUDA IPC */
        LK_CUDA(e, cudaMemset(e->d_mbox, 0, sizeof(PeerMbox)));
        LK_CUDA(e, cudaDeviceSynchronize());
        e->peer_seq = 0;
    }
    if (handle) {
        cudaIpcMemHandle_t h;
        static_assert(sizeof(h) == LKCUDA_PEER_HANDLE_BYTES, "CUDA IPC handle size");
        LK_CUDA(e, cudaIpcGetMemHandle(&h, e->d_mbox));
        memcpy(handle, &h, sizeof(h));
    }
    if (local_ptr) *local_ptr = e->d_mbox;
    return LKCUDA_OK;
}

int lkcuda_detach_peers(lkcuda_engine *e) {
    LK_ENTER(e);
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    e->v.peer_world = 1; e->v.peer_rank = 0; e->v.peer_mbox = NULL;
    for (int q = 0; q < LKCUDA_MAX_PEERS; q++)
        if (e->peer_opened[q]) { cudaIpcCloseMemHandle(e->peer_opened[q]); e->peer_opened[q] = NULL; }
    return LKCUDA_OK;
}

int lkcuda_attach_peers(lkcuda_engine *e, int32_t rank, int32_t world, const uint8_t *handles, void *const *raw_ptrs) {
    LK_ENTER(e);
    if (world < 1 || world > LKCUDA_MAX_PEERS || rank < 0 || rank >= world) return fail(e, LKCUDA_ERR_ARG, "lkcuda_attach_peers: bad rank / world (at most %d peers)", LKCUDA_MAX_PEERS);
    if (!e->d_mbox) return fail(e, LKCUDA_ERR_STATE, "lkcuda_attach_peers: call lkcuda_peer_handle first");
    if (world > 1 && !handles && !raw_ptrs) return fail(e, LKCUDA_ERR_ARG, "lkcuda_attach_peers: no peer addresses");
    if (e->peer_seq != 0) return fail(e, LKCUDA_ERR_STATE, "lkcuda_attach_peers: the mailbox has already been used");
    int rc0 = lkcuda_detach_peers(e);
    if (rc0) return rc0;
    PeerMbox *tab[LKCUDA_MAX_PEERS];
    memset(tab, 0, sizeof(tab));
    for (int q = 0; q < world; q++) {
        if (q == rank) { tab[q] = e->d_mbox; continue; }
        if (raw_ptrs && raw_ptrs[q]) {
            cudaPointerAttributes at;
            if (cudaPointerGetAttributes(&at, raw_ptrs[q]) != cudaSuccess || at.type != cudaMemoryTypeDevice) {
                cudaGetLastError();
                return fail(e, LKCUDA_ERR_ARG, "lkcuda_attach_peers: peer %d is not device memory", q);
            }
            if (at.device != e->device) {
                int can = 0;
                cudaDeviceCanAccessPeer(&can, e->device, at.device);
                if (!can) return fail(e, LKCUDA_ERR_STATE, "lkcuda_attach_peers: device %d cannot access device %d", e->device, at.device);
                cudaError_t pe = cudaDeviceEnablePeerAccess(at.device, 0);
                if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(e, pe, "cudaDeviceEnablePeerAccess");
                cudaGetLastError();
            }
            tab[q] = (PeerMbox *)raw_ptrs[q];
        } else if (handles) {
            cudaIpcMemHandle_t h;
            memcpy(&h, handles + (size_t)q * LKCUDA_PEER_HANDLE_BYTES, sizeof(h));
            void *p = NULL;
            cudaError_t ie = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
            if (ie != cudaSuccess) { cudaGetLastError(); lkcuda_detach_peers(e); return cuda_fail(e, ie, "cudaIpcOpenMemHandle (peer mailbox)"); }
            e->peer_opened[q] = p;
            tab[q] = (PeerMbox *)p;
        } else return fail(e, LKCUDA_ERR_ARG, "lkcuda_attach_peers: no address for peer %d", q);
    }
    if (!e->d_peer_tab) LK_CUDA(e, cudaMalloc((void **)&e->d_peer_tab, sizeof(tab)));
    LK_CUDA(e, cudaMemcpy(e->d_peer_tab, tab, sizeof(tab), cudaMemcpyHostToDevice));
    if (world > 1) { preload_reduce_kernels(); e->v.peer_world = world; e->v.peer_rank = rank; e->v.peer_mbox = e->d_peer_tab; }
    return LKCUDA_OK;
}

int lkcuda_result_device_ptr(lkcuda_engine *e, void **p) {
    if (!e || !p) return LKCUDA_ERR_ARG;
    *p = e->v.result;
    return LKCUDA_OK;
}

int lkcuda_set_result_buffer(lkcuda_engine *e, void *dev_ptr) {
    LK_ENTER(e);
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    e->v.result = dev_ptr ? (double *)dev_ptr : e->d_result;
    return LKCUDA_OK;
}

int lkcuda_fetch_pattern_lh(lkcuda_engine *e, double *dst) {
    LK_ENTER(e);
    if (!dst) return fail(e, LKCUDA_ERR_ARG, "null destination");
    if (e->v.P) LK_CUDA(e, cudaMemcpyAsync(dst, e->d_plh, (size_t)e->v.P * 8, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    return upload_verdict(e);
}

int lkcuda_fetch_pattern_lh_cat(lkcuda_engine *e, double *dst) {
    LK_ENTER(e);
    if (!dst) return fail(e, LKCUDA_ERR_ARG, "null destination");
    if (e->v.P) LK_CUDA(e, cudaMemcpyAsync(dst, e->d_plhcat, (size_t)e->v.P * e->v.K * 8, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    return LKCUDA_OK;
}

int lkcuda_fetch_vector(lkcuda_engine *e, int32_t vec, double *partial_lh, int16_t *scale_num) {
    LK_ENTER(e);
    if (vec < 0 || vec >= e->cfg.nvec) return fail(e, LKCUDA_ERR_ARG, "vector index out of range");
    const DevView &v = e->v;
    if (partial_lh && v.NP) LK_CUDA(e, cudaMemcpyAsync(partial_lh, e->v.arena + (size_t)vec * v.vec_stride, (size_t)v.NP * v.B * 8, cudaMemcpyDeviceToHost, e->stream));
    if (scale_num && v.NP) LK_CUDA(e, cudaMemcpyAsync(scale_num, e->v.scale + (size_t)vec * v.sc_stride, (size_t)v.NP * 2, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    return LKCUDA_OK;
}

int lkcuda_store_vector(lkcuda_engine *e, int32_t vec, const double *partial_lh, const int16_t *scale_num) {
    LK_ENTER(e);
    if (vec < 0 || vec >= e->cfg.nvec) return fail(e, LKCUDA_ERR_ARG, "vector index out of range");
    const DevView &v = e->v;
    if (partial_lh && v.NP) LK_CUDA(e, cudaMemcpyAsync(e->v.arena + (size_t)vec * v.vec_stride, partial_lh, (size_t)v.NP * v.B * 8, cudaMemcpyHostToDevice, e->stream));
    if (scale_num && v.NP) LK_CUDA(e, cudaMemcpyAsync(e->v.scale + (size_t)vec * v.sc_stride, scale_num, (size_t)v.NP * 2, cudaMemcpyHostToDevice, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    e->theta_ready = false;
    return LKCUDA_OK;
}

int lkcuda_fetch_theta(lkcuda_engine *e, double *theta) {
    LK_ENTER(e);
    if (!theta) return fail(e, LKCUDA_ERR_ARG, "null destination");
    if (e->v.NP) LK_CUDA(e, cudaMemcpyAsync(theta, e->d_theta, (size_t)e->v.NP * e->v.B * 8, cudaMemcpyDeviceToHost, e->stream));
    LK_CUDA(e, cudaStreamSynchronize(e->stream));
    return LKCUDA_OK;
}

int lkcuda_vector_device_ptr(lkcuda_engine *e, int32_t vec, void **partial_lh, void **scale_num) {
    if (!e || vec < 0 || vec >= e->cfg.nvec) return LKCUDA_ERR_ARG;
    if (partial_lh) *partial_lh = e->v.arena + (size_t)vec * e->v.vec_stride;
    if (scale_num) *scale_num = e->v.scale + (size_t)vec * e->v.sc_stride;
    return LKCUDA_OK;
}

int lkcuda_launch_count(const lkcuda_engine *e, uint64_t *c) { if (!e || !c) return LKCUDA_ERR_ARG; *c = e->launches; return LKCUDA_OK; }
int lkcuda_algorithmic_bytes(const lkcuda_engine *e, uint64_t *b) { if (!e || !b) return LKCUDA_ERR_ARG; *b = e->bytes; return LKCUDA_OK; }

int lkcuda_reset_counters(lkcuda_engine *e) {
    LK_ENTER(e);
    drain_events(e);
    e->launches = 0; e->bytes = 0;
    memset(e->kind_ms, 0, sizeof(e->kind_ms)); memset(e->kind_launches, 0, sizeof(e->kind_launches)); memset(e->kind_bytes, 0, sizeof(e->kind_bytes)); memset(e->kind_flops, 0, sizeof(e->kind_flops));
    return LKCUDA_OK;
}

int lkcuda_set_profiling(lkcuda_engine *e, int32_t on) {
    LK_ENTER(e);
    drain_events(e);
    e->profiling = on != 0;
    return LKCUDA_OK;
}

int lkcuda_kernel_flops(lkcuda_engine *e, int32_t kind, uint64_t *flops) {
    if (!e || !flops || kind < 0 || kind >= KIND_COUNT) return LKCUDA_ERR_ARG;
    *flops = e->kind_flops[kind];
    return LKCUDA_OK;
}

int lkcuda_kernel_stats(lkcuda_engine *e, int32_t kind, uint64_t *launches, uint64_t *bytes, double *device_ms) {
    LK_ENTER(e);
    if (kind < 0 || kind >= KIND_COUNT) return fail(e, LKCUDA_ERR_ARG, "bad kernel kind");
    drain_events(e);
    if (launches) *launches = e->kind_launches[kind];
    if (bytes) *bytes = e->kind_bytes[kind];
    if (device_ms) *device_ms = e->kind_ms[kind];
    return LKCUDA_OK;
}

}  // extern "C"
