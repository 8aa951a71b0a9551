/*
 * lkcuda_common.cuh — shared declarations of the sm_100a likelihood engine behind include/lkcuda.h.
 *
 * Data layout in HBM (one engine = one pattern shard on one device), NP = P + U patterns:
 *   arena     double[nvec][vec_stride]      vector v = partial_lh[ptn][blk][state], the reference's layout
 *                                           (phylotree.cpp:863-974); vec_stride = NP*B rounded up to 256 B
 *   scale     int16 [nvec][sc_stride]       scale_num[ptn] (UBYTE = short, phylonode.h:17)
 *   states    uint8 [ntaxa][st_stride]      taxon-major tip states (pattern-major strings in the reference)
 *   theta     double[NP*B]                  theta_all (phylotree.cpp:687)
 *   pattern_lh double[NP], pattern_lh_cat double[P*K], ptn_freq double[P], ptn_invar double[NP]
 * Every vector starts on a 256-byte boundary and B*8 is normally a multiple of 32, so a lane always
 * moves whole 32-byte sectors (LDG.E.256 / STG.E.256) whatever the thread<->pattern mapping is.
 */
#ifndef LKCUDA_COMMON_CUH
#define LKCUDA_COMMON_CUH

#include "../../include/lkcuda.h"
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <vector>

namespace lkc {

/* phylotree.h:52-54 */
#define LK_THR_INV 0x1p256
#define LK_THR 0x1p-256
#define LK_LOG_THR (-177.44567822334599)   /* log(2^-256) */

enum KernelKind {
    KIND_PREP = 0, KIND_PARTIAL = 1, KIND_BRANCH = 2, KIND_THETA = 3, KIND_DERV = 4, KIND_FROMBUF = 5,
    KIND_NEWTON = 6, KIND_MISC = 7, KIND_RELL = 8, KIND_EM = 9, KIND_COUNT = 10
};

/* cudaFuncAttributeMaxDynamicSharedMemorySize is a property of the FUNCTION (per context), not of an engine: every engine sets the
   same value - the device maximum - so that engines of different shapes, driven from different host threads, never lower it
   under each other's launches */
#define LK_MAX_DYN_SMEM (227 * 1024)

/* slots of the cross-block reduction buffer */
enum { RED_SLOTS = 8 };

struct PendingEvent { int kind; cudaEvent_t start, stop; };

/* device-side op record of one node of the schedule */
struct DevOp {
    int32_t dst;        /* destination vector                                   */
    int32_t nchild;     /* 2 for a bifurcating node                             */
    int32_t child0;     /* first entry in the child list / child-table slots    */
    int32_t pad;
};

struct DevChild {
    int32_t id;         /* vector (>=0) or LKCUDA_TIP(taxon)                    */
    int32_t pad;        /* 0, or 1 + the device length slot that holds this branch's length (lkcuda.h: length slots) */
    double  len;
};

struct PeerMbox;

/* B = 16 (S = 4, K = 4): everything the reductions need to form exp(eval * rate * len) * weight, as a KERNEL PARAMETER - the lanes build
   their table entries without a dependent chain of global loads and without a block barrier */
struct B16Tab { double cof[16]; double w[4]; int eig[4]; };

/* everything a kernel needs to know about the engine, passed by value */
struct DevView {
    int S, K, E, B;             /* states, blocks per pattern, eigensystems, B = S*K        */
    int SU;                     /* state_unknown                                             */
    int flags;
    long long P, U, NP;         /* observed, unobserved, total patterns                      */
    long long vec_stride, sc_stride, st_stride;
    double nsite;
    const uint8_t *states;
    const double *freq, *invar;
    const double *eval, *evec, *inv_evec, *blk_rate, *blk_weight, *tip;
    const int32_t *blk_eigen;
    double *arena; int16_t *scale;
    double *theta; int16_t *theta_sc; double *theta_meta;   /* theta_meta[0] = scale sum of the theta branch */
    double *pattern_lh, *pattern_lh_cat;
    uint4 *red; unsigned int *counter; double *result;     /* red: 16-byte slots of the cross-block reduction (grid_reduce) */
    /* per-child transition tables built by the prep kernel:
       slot c: ech[K*S*S] then tab[(SU+1)*B]                                                   */
    double *tables; long long table_slot;   /* doubles per slot */
    /* result publication (publish_result): the host-mapped copy of the result block + the sequence number of this launch, and the
       peer mailboxes of the fused cross-GPU all-reduce (lkcuda_attach_peers) */
    double *host_result; unsigned int seq;
    int peer_world, peer_rank; unsigned int peer_seq;
    PeerMbox *const *peer_mbox;
    /* device length slots (lkcuda.h): lens[LKCUDA_NLENSLOT] branch lengths written by the persistent Newton kernels and read by the table
       kernels / lnL-from-buffer of later launches, then [LKCUDA_NLENSLOT][4] = {optx, d2l, evaluations, status} of the run that wrote a slot */
    double *lens;
    /* site-specific frequency models (LKCUDA_FLAG_SITEMODEL): pattern -> eigensystem, tip code -> set of plain states */
    const int32_t *ptn_model; const unsigned long long *state_mask;
};

/* one per engine, in cudaMalloc'ed (IPC-exportable) device memory: rank r writes its 8 sums into data[parity][r] of EVERY rank's
   mailbox, then flag[parity][r] = sequence number; the owner adds the deposits in rank order once all flags show the number */
struct PeerMbox {
    double data[2][LKCUDA_MAX_PEERS][LKCUDA_NRESULT];
    unsigned int flag[2][LKCUDA_MAX_PEERS];
};

}  // namespace lkc

struct lkcuda_aux;      /* lkcuda_aux.cu: RELL matrix and _pattern_lh_cat work buffers (allocated on first use) */

struct lkcuda_engine {
    lkcuda_config cfg;
    lkcuda_aux *aux;
    lkc::DevView v;
    int device;
    cudaStream_t stream; bool own_stream;
    /* owned device allocations */
    uint8_t *d_states; double *d_freq, *d_invar;
    double *d_eval, *d_evec, *d_inv, *d_rate, *d_weight, *d_tip; int32_t *d_eigen;
    double *d_arena; int16_t *d_scale;      /* the engine's OWN arena; v.arena / v.scale may point at a caller-owned one (lkcuda_attach_arena) */
    long long own_vec_stride, own_sc_stride; int own_nvec;
    double *d_theta; int16_t *d_theta_sc; double *d_theta_meta;
    int32_t *d_ptn_model; unsigned long long *d_state_mask;     /* LKCUDA_FLAG_SITEMODEL */
    double *d_plh, *d_plhcat, *d_plh_out;   /* d_plh_out: scale-corrected copy of _pattern_lh on its way to the host (lazy) */
    cudaStream_t dstream; cudaEvent_t ev_plh_ready, ev_plh_done; bool plh_pending;   /* asynchronous device-to-host fetch of d_plh_out (lazy) */
    uint4 *d_red; unsigned int *d_counter; double *d_result;
    double *h_result;                       /* pinned */
    double *h_mapped;                       /* pinned + mapped: 8 slots {lo32, tag, hi32, tag}, tag = sequence number | error << 31, written by publish_result */
    lkc::B16Tab b16tab;                     /* DNA + 4 categories: eval * rate, weights and eigensystem of the 4 blocks, passed to the B = 16 kernels by value */
    unsigned int result_seq;                /* sequence number of the last publishing launch */
    lkc::PeerMbox *d_mbox; lkc::PeerMbox **d_peer_tab; void *peer_opened[LKCUDA_MAX_PEERS]; unsigned int peer_seq;
    double *d_newton_out; double *h_newton_out;
    double *d_lens; double *h_lens;         /* device length slots and the pinned buffer lkcuda_len_slots_fetch copies them to */
    void *d_newton_ctl; int newton_max_grid;   /* state of the persistent Newton kernel; the largest grid that is co-resident */
    /* transition tables */
    double *d_tables; long long table_slots_cap;
    /* op upload ring (pinned host + device) */
    enum { RING = 4 };
    lkc::DevOp *h_ops[RING], *d_ops[RING]; lkc::DevChild *h_child[RING], *d_child[RING];
    int ring_cap_ops[RING], ring_cap_child[RING]; cudaEvent_t ring_ev[RING]; bool ring_used[RING]; int ring_next;
    int red_blocks;                         /* capacity of d_red in blocks */
    size_t trav_smem_set;                   /* dynamic shared memory the traversal kernel is currently allowed */
    double *d_inv_frags;                    /* fragment-ordered inverse eigenvectors of the protein MMA kernel */
    size_t mma_smem_set; bool trav_disable_mma;   /* LKCUDA_NO_MMA=1: protein falls back to the generic kernel (test knob) */
    int trav_maxslots;                      /* LKCUDA_TRV_MAXSLOTS: cap on shared-memory slots (-1 = what fits) */
    std::vector<char> trav_mark;            /* launch_traverse: vectors written by the list being scheduled (all zero between calls) */
    /* pipelined tip-state upload (lkcuda_set_tip_states_async): chunks of patterns go up on `cstream`; after each chunk the copy
       engine bumps *d_upload_flag, which the DNA traversal kernel's producer warp polls before it stages a tile's tip states */
    enum { UPLOAD_MAX_CHUNKS = 64 };
    cudaStream_t cstream; cudaEvent_t ev_upload_done, ev_main_mark;
    unsigned int *d_upload_flag, *h_upload_vals, *h_upload_bad;    /* device counter; pinned sources of its values; pinned validation result */
    unsigned int upload_base; int upload_chunks; long long upload_chunk_pat;
    bool upload_pending;        /* the main stream has not yet been ordered after the upload */
    bool upload_check;          /* the range check of the uploaded states has not been examined yet */
    bool theta_ready; int theta_dad; int theta_node;
    /* lkcuda_partial_batch_lnl: root branch the LAST piece of the current op list may evaluate in the tail of its traversal launch
       (fuse_state: 0 = no request, 1 = requested, 2 = the last piece is being issued, 3 = done by the traversal kernel) */
    int fuse_state, fuse_dad, fuse_node; double fuse_len;
    bool model_set, states_set, weights_set;
    int num_sms;
    uint64_t launches, bytes;
    /* optional per-kernel-kind timing */
    bool profiling;
    std::vector<lkc::PendingEvent> pending;
    std::vector<cudaEvent_t> event_pool;
    double kind_ms[lkc::KIND_COUNT]; uint64_t kind_launches[lkc::KIND_COUNT]; uint64_t kind_bytes[lkc::KIND_COUNT]; uint64_t kind_flops[lkc::KIND_COUNT];
    uint64_t total_dev_bytes;
    char err[512];
};

namespace lkc {

int fail(lkcuda_engine *e, int code, const char *fmt, ...);
int cuda_fail(lkcuda_engine *e, cudaError_t rc, const char *what);

#define LK_CUDA(e, call) do { cudaError_t rc_ = (call); if (rc_ != cudaSuccess) return lkc::cuda_fail((e), rc_, #call); } while (0)
#define LK_ENTER(e) do { if (!(e)) return LKCUDA_ERR_ARG; cudaError_t rc_ = cudaSetDevice((e)->device); \
                         if (rc_ != cudaSuccess) return lkc::cuda_fail((e), rc_, "cudaSetDevice"); } while (0)

/* bracket a launch for statistics (and CUDA-event timing when profiling is on) */
void launch_begin(lkcuda_engine *e, int kind);
int launch_end(lkcuda_engine *e, int kind, uint64_t algorithmic_bytes, const char *what);
void drain_events(lkcuda_engine *e);
void aux_destroy(lkcuda_engine *e);
void preload_reduce_kernels();      /* lkcuda_reduce.cu: force-load every reduction kernel (called when peers are attached) */

inline bool valid_end(const lkcuda_engine *e, int32_t id) {
    return id >= 0 ? id < e->cfg.nvec : LKCUDA_TIP_TAXON(id) < e->cfg.ntaxa;
}

/* call right before launching a kernel that ends in publish_result(): stamps e->v with the launch's sequence numbers */
inline void next_result(lkcuda_engine *e) {
    e->result_seq = (e->result_seq + 1u) & 0x7fffffffu;      /* 31 bits travel in the tags of the host-mapped slots */
    if (!e->result_seq) e->result_seq = 1u;
    e->v.seq = e->result_seq;
    if (e->v.peer_world > 1) e->v.peer_seq = ++e->peer_seq;
}

/* implemented in lkcuda_partial.cu / lkcuda_reduce.cu */
int ensure_tables(lkcuda_engine *e, long long slots);
int branch_lnl_followups(lkcuda_engine *e, int32_t dad, int32_t node, double len, double *result);   /* lkcuda_reduce.cu */
/* order the engine stream after a pipelined tip-state upload (every kernel that reads tip states without polling calls this) */
int states_barrier(lkcuda_engine *e);
/* after a synchronisation point: report a failed range check of an asynchronous upload */
int upload_verdict(lkcuda_engine *e);

}  // namespace lkc

/* ------------------------------------------------------------------ device helpers ------------------------------------------------------------------ */
#ifdef __CUDACC__
namespace lkc {

struct __align__(32) d4 { double x, y, z, w; };

/* length of the branch to a child: the value the caller gave, or the device slot a Newton run of an earlier launch left it in */
__device__ __forceinline__ double child_len(const DevView &v, const DevChild &ch) {
    return ch.pad ? __ldcg(v.lens + (ch.pad - 1)) : ch.len;
}

/* whole-sector streaming accesses: LDG.E.256 / STG.E.256 on sm_100a */
__device__ __forceinline__ d4 ld_stream256(const double *p) {
    d4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream256(double *p, const d4 &v) {
    asm volatile("st.global.L1::no_allocate.v4.f64 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "d"(v.x), "d"(v.y), "d"(v.z), "d"(v.w) : "memory");
}

__device__ __forceinline__ const double *vec_lh(const DevView &v, int id) { return v.arena + (long long)id * v.vec_stride; }
__device__ __forceinline__ double *vec_lh_w(const DevView &v, int id) { return v.arena + (long long)id * v.vec_stride; }
__device__ __forceinline__ const int16_t *vec_sc(const DevView &v, int id) { return v.scale + (long long)id * v.sc_stride; }
__device__ __forceinline__ int16_t *vec_sc_w(const DevView &v, int id) { return v.scale + (long long)id * v.sc_stride; }
__device__ __forceinline__ const uint8_t *tip_states(const DevView &v, int id) { return v.states + (long long)LKCUDA_TIP_TAXON(id) * v.st_stride; }

/* fixed-order block reduction of N doubles per thread; result valid in thread 0.
   `sh` must hold N * 32 doubles. Deterministic for a fixed block size. */
template <int N>
__device__ __forceinline__ void block_reduce(double (&val)[N], double *sh) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int i = 0; i < N; i++) {
        double x = val[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
        if (lane == 0) sh[i * 32 + warp] = x;
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int i = 0; i < N; i++) {
            double x = lane < nwarp ? sh[i * 32 + lane] : 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) if (o < nwarp) x += __shfl_down_sync(0xffffffffu, x, o);    /* lanes >= nwarp hold zeros: the skipped levels would add nothing */
            val[i] = x;
        }
    }
}

/* cross-block stage WITHOUT fences and atomics.  Thread 0 of every block deposits its N sums as 16-byte slots {lo32, tag, hi32, tag}:
   each 8-byte unit carries its own tag (the scheme of NCCL's LL protocol; 8-byte stores are single transactions), tag = generation of
   this reduction = *counter + 1.  Block 0 is the finisher: after its own deposit its threads pick up the blocks' slots (spinning on a slot
   whose tags do not show the generation yet), add them in BLOCK order and return true in thread 0 with the totals in `val`; it then
   advances *counter.  Launches of one engine are serialised on its stream, and the persistent Newton kernel only starts the next
   reduction after the finisher released the next step, so a generation is never written while the previous one is still being read.
   Deterministic run to run (fixed summation tree).  Measured on B200 (tools/reduce_bench.py, r2): the store + __threadfence + ticket
   atomic + fence + re-read chain of the round-1 version cost ~3000 cycles after the last block's stream; this one ~1500. */
template <int N>
__device__ __forceinline__ bool grid_reduce(double (&val)[N], double *sh, uint4 *red, unsigned int *counter) {
    unsigned int gen = 0;
    if (threadIdx.x == 0 || blockIdx.x == 0) { gen = __ldcg(counter) + 1u; if (!gen) gen = 1u; }      /* issued before the block reduction: off the critical path */
    block_reduce<N>(val, sh);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < N; i++) {
            const unsigned long long bits = (unsigned long long)__double_as_longlong(val[i]);
            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(red + (size_t)blockIdx.x * RED_SLOTS + i),
                         "r"((unsigned int)bits), "r"(gen), "r"((unsigned int)(bits >> 32)), "r"(gen) : "memory");
        }
    }
    if (blockIdx.x != 0) return false;
    __syncthreads();        /* sh is reused below */
    double acc[N];
#pragma unroll
    for (int i = 0; i < N; i++) acc[i] = 0.0;
    const long long t0 = clock64();
    bool lost = false;
    for (unsigned int b = threadIdx.x; b < gridDim.x; b += blockDim.x) {
        const uint4 *src = red + (size_t)b * RED_SLOTS;
        uint4 q[N];
        for (;;) {
            bool ok = true;
#pragma unroll
            for (int i = 0; i < N; i++) {
                asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(q[i].x), "=r"(q[i].y), "=r"(q[i].z), "=r"(q[i].w) : "l"(src + i) : "memory");
                ok = ok && q[i].y == gen && q[i].w == gen;
            }
            if (ok) break;
            if (clock64() - t0 > 8000000000ll) { lost = true; break; }     /* ~4 s: a block never deposited (cannot happen in a healthy launch); do not hang the GPU */
        }
#pragma unroll
        for (int i = 0; i < N; i++) acc[i] += lost ? __longlong_as_double(0x7ff8000000000000ll) : __longlong_as_double((long long)(((unsigned long long)q[i].z << 32) | q[i].x));
    }
    block_reduce<N>(acc, sh);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < N; i++) val[i] = acc[i];
        *reinterpret_cast<volatile unsigned int *>(counter) = gen;
    }
    return threadIdx.x == 0;
}

/* ---- publication of a reduction's result block (called by ONE thread: the one grid_reduce returned true in) ----
   1. fused cross-GPU all-reduce (only the scalar sums cross NVLink): the block goes straight into every peer's mailbox with
      plain P2P stores, a release-store of the sequence number follows; the thread then waits for all ranks' numbers in its OWN mailbox
      and adds the deposits in rank order - every rank gets bit-identical sums, no NCCL launch, no host round trip.  Slot 6
      (log(1 - prob_const), read by asc_adjust_kernel) is replicated, not additive: it stays local.
   2. the block is written to device memory (v.result) and to host-mapped memory followed by the sequence number, which
      lkcuda_wait_result() polls instead of issuing a copy and a stream synchronisation. */
__device__ __forceinline__ unsigned long long lk_globaltimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

/* the exchange alone (step 1 above) under sequence number `seq`; returns 1 when a peer did not show up in time */
static __device__ __noinline__ unsigned int peer_exchange(const DevView &v, double (&r)[LKCUDA_NRESULT], unsigned int seq) {
    unsigned int err = 0;
    const int par = (int)(seq & 1u);
    for (int q = 0; q < v.peer_world; q++) {
        volatile double *dst = v.peer_mbox[q]->data[par][v.peer_rank];
#pragma unroll
        for (int i = 0; i < LKCUDA_NRESULT; i++) dst[i] = r[i];
    }
    __threadfence_system();
    for (int q = 0; q < v.peer_world; q++) {
        unsigned int *f = &v.peer_mbox[q]->flag[par][v.peer_rank];
        asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(f), "r"(seq) : "memory");
    }
    PeerMbox *me = v.peer_mbox[v.peer_rank];
    const unsigned long long t0 = lk_globaltimer();
    for (int q = 0; q < v.peer_world && !err; q++) {
        const unsigned int *f = &me->flag[par][q];
        unsigned int seen;
        for (;;) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(f) : "memory");
            if (seen == seq) break;
            if (lk_globaltimer() - t0 > 10000000000ull) { err = 1; break; }      /* 10 s: a peer never launched its reduction */
        }
    }
    if (!err) {
        const double own6 = r[6];
#pragma unroll
        for (int i = 0; i < LKCUDA_NRESULT; i++) r[i] = 0.0;
        for (int q = 0; q < v.peer_world; q++) {
            const volatile double *src = me->data[par][q];
#pragma unroll
            for (int i = 0; i < LKCUDA_NRESULT; i++) r[i] += src[i];
        }
        r[6] = own6;
    }
    return err;
}

/* steps 2 and 3: device result block, host-mapped slots {lo32, tag, hi32, tag} */
static __device__ __forceinline__ void post_result(const DevView &v, const double (&r)[LKCUDA_NRESULT], unsigned int err) {
#pragma unroll
    for (int i = 0; i < LKCUDA_NRESULT; i++) v.result[i] = r[i];
    if (v.host_result) {
        /* slot i = {lo32, tag, hi32, tag}: every 8-byte unit carries the sequence number (| error << 31), so the host can accept a value
           as soon as both of its tags show the number - no fence.sys (a PCIe round trip) between data and flag */
        const unsigned int tag = (v.seq & 0x7fffffffu) | (err << 31);
#pragma unroll
        for (int i = 0; i < LKCUDA_NRESULT; i++) {
            const unsigned long long bits = (unsigned long long)__double_as_longlong(r[i]);
            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(v.host_result + 2 * i),
                         "r"((unsigned int)bits), "r"(tag), "r"((unsigned int)(bits >> 32)), "r"(tag) : "memory");
        }
    }
}

static __device__ __forceinline__ void publish_result(const DevView &v, double (&r)[LKCUDA_NRESULT]) {
    unsigned int err = 0;
    if (v.peer_world > 1) err = peer_exchange(v, r, v.peer_seq);
    post_result(v, r, err);
}

/* ---- the same two stages for PART of a block: threads 0 .. nthr-1 (whole warps) of a kernel whose other warps have other roles
   (the traversal kernel's consumer warps).  Synchronises on named barrier 1 instead of __syncthreads(). ---- */
__device__ __forceinline__ void bar_part(int nthr) { asm volatile("bar.sync 1, %0;" :: "r"(nthr) : "memory"); }

template <int N>
__device__ __forceinline__ void block_reduce_part(double (&val)[N], double *sh, int nthr) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = nthr >> 5;
#pragma unroll
    for (int i = 0; i < N; i++) {
        double x = val[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
        if (lane == 0) sh[i * 32 + warp] = x;
    }
    bar_part(nthr);
    if (warp == 0) {
#pragma unroll
        for (int i = 0; i < N; i++) {
            double x = lane < nwarp ? sh[i * 32 + lane] : 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
            val[i] = x;
        }
    }
}

template <int N>
__device__ __forceinline__ bool grid_reduce_part(double (&val)[N], double *sh, uint4 *red, unsigned int *counter, int nthr) {
    unsigned int gen = 0;
    if (threadIdx.x == 0 || blockIdx.x == 0) { gen = __ldcg(counter) + 1u; if (!gen) gen = 1u; }
    block_reduce_part<N>(val, sh, nthr);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < N; i++) {
            const unsigned long long bits = (unsigned long long)__double_as_longlong(val[i]);
            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(red + (size_t)blockIdx.x * RED_SLOTS + i),
                         "r"((unsigned int)bits), "r"(gen), "r"((unsigned int)(bits >> 32)), "r"(gen) : "memory");
        }
    }
    if (blockIdx.x != 0) return false;
    bar_part(nthr);
    double acc[N];
#pragma unroll
    for (int i = 0; i < N; i++) acc[i] = 0.0;
    const long long t0 = clock64();
    bool lost = false;
    for (unsigned int b = threadIdx.x; b < gridDim.x; b += (unsigned int)nthr) {
        const uint4 *src = red + (size_t)b * RED_SLOTS;
        uint4 q[N];
        for (;;) {
            bool ok = true;
#pragma unroll
            for (int i = 0; i < N; i++) {
                asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(q[i].x), "=r"(q[i].y), "=r"(q[i].z), "=r"(q[i].w) : "l"(src + i) : "memory");
                ok = ok && q[i].y == gen && q[i].w == gen;
            }
            if (ok) break;
            if (clock64() - t0 > 8000000000ll) { lost = true; break; }
        }
#pragma unroll
        for (int i = 0; i < N; i++) acc[i] += lost ? __longlong_as_double(0x7ff8000000000000ll) : __longlong_as_double((long long)(((unsigned long long)q[i].z << 32) | q[i].x));
    }
    block_reduce_part<N>(acc, sh, nthr);
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < N; i++) val[i] = acc[i];
        *reinterpret_cast<volatile unsigned int *>(counter) = gen;
    }
    return threadIdx.x == 0;
}

/* log, NaN repair, pattern_lh, frequency weighting, +ASC accumulation: the common tail of branch lnL and
   lnL-from-buffer (phylokernel.h:828-866, 895-900, 1009-1016). acc[0..3] = {sum freq*log, scale sum, prob_const, repaired} */
__device__ __forceinline__ void lnl_tail_vals(const DevView &v, long long p, double lh, int sc, double inv, double f, double (&acc)[4]) {
    if (p < v.P) {
        double l = log(fabs(lh + inv));
        if (isnan(l) || isinf(l)) { l = LK_LOG_THR * 4.0; acc[3] += 1.0; }
        v.pattern_lh[p] = l;
        acc[0] += l * f;
        if (sc) acc[1] += LK_LOG_THR * f * (double)sc;
    } else {
        if (sc >= 1) lh *= LK_THR;
        acc[2] += lh + inv;
    }
}
__device__ __forceinline__ void lnl_tail(const DevView &v, long long p, double lh, int sc, double (&acc)[4]) {
    lnl_tail_vals(v, p, lh, sc, v.invar[p], p < v.P ? v.freq[p] : 0.0, acc);
}

__device__ __forceinline__ void lnl_finish(const DevView &v, const double (&acc)[4], double scale_sum) {
    double lnl = acc[0] + scale_sum, lc = 0.0;
    if (v.U > 0) { lc = log(1.0 - acc[2]); lnl -= v.nsite * lc; }
    double r[LKCUDA_NRESULT] = {lnl, 0.0, 0.0, acc[2], acc[3], scale_sum, lc, 0.0};
    publish_result(v, r);
}

}  // namespace lkc
#endif  /* __CUDACC__ */
#endif
