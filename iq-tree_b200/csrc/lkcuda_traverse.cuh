/*
 * lkcuda_traverse.cuh — pieces shared by the whole-traversal kernels (lkcuda_partial.cu: DNA, FP64 FMA;
 * lkcuda_partial_mma.cu: protein, FP64 tensor-core MMA): the node descriptor the producer warp streams, the
 * mbarrier / bulk-copy primitives of the shared-memory ring, and the host-side schedule (order + slot allocation).
 */
#ifndef LKCUDA_TRAVERSE_CUH
#define LKCUDA_TRAVERSE_CUH

#include "lkcuda_common.cuh"

#define TRV_SLOT_PREV (-3)  /* child 1 is the result of the node executed immediately before: still in registers */

namespace lkc {

/* node descriptor, written by the host, streamed into the ring by the producer (64 bytes) */
struct TravOp {
    double *dst_lh; int16_t *dst_sc;    /* arena vector (partial_lh, scale_num) receiving the result */
    const double *src0_lh, *src1_lh;    /* arena vectors of the children (NULL for a tip) */
    int32_t src0, src1;                 /* child: vector (>= 0) or LKCUDA_TIP(taxon) */
    int32_t slot0, slot1;               /* shared-memory slot holding the child (-1: read the arena vector; slot1 may be TRV_SLOT_PREV) */
    int32_t dst_slot;                   /* slot the result is parked in (-1: not needed again in this launch or consumed by the next
                                           node from registers, -2: needed but no slot was free: the consumer re-reads the arena) */
    int32_t dst;
    int32_t pad[2];
};

#ifdef __CUDACC__
__device__ __forceinline__ d4 ld_coherent256(const double *p) { /* may have been written earlier by this very thread */
    d4 r;
    asm volatile("ld.global.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ double2 ld_coherent128(const double *p) {
    double2 r;
    asm volatile("ld.global.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity), "r"(0x989680u) : "memory");   /* suspend-time hint: the warp sleeps in hardware instead of spinning on issue slots */
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) { while (!mbar_try(bar, parity)) { } }
__device__ __forceinline__ void mbar_wait_backoff(uint64_t *bar, uint32_t parity) { while (!mbar_try(bar, parity)) __nanosleep(256); }
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

#endif  /* __CUDACC__ */

struct TravNode { int op; int32_t src0, slot0, src1, slot1, dst_slot; double len0, len1; int32_t ref0, ref1; /* 1 + length slot, or 0 */ };
struct TravSchedule { std::vector<TravNode> nodes; int nslots; };

/* implemented in lkcuda_partial.cu */
int schedule_traverse(lkcuda_engine *e, const lkcuda_partial_op *ops, int nops, int maxslots, bool use_prev, TravSchedule &sch);
int ring_acquire(lkcuda_engine *e, int nops, int nchild, int *slot_out);
uint64_t op_bytes(const DevView &v, int ntip, int nint);
uint64_t op_flops(const DevView &v, int ntip, int nint);

/* implemented in lkcuda_partial_mma.cu */
bool use_traverse_mma(const lkcuda_engine *e);
int launch_traverse_mma(lkcuda_engine *e, const lkcuda_partial_op *ops, int nops);

/* implemented in lkcuda_partial_mmak.cu (codon, wide-block protein) */
long long mmak_table_slot(const DevView &v);
bool use_traverse_mmak(const lkcuda_engine *e);
int launch_traverse_mmak(lkcuda_engine *e, const lkcuda_partial_op *ops, int nops);

/* implemented in lkcuda_sitemodel.cu (site-specific frequency models): ops / children already written to ring slot `slot` */
int launch_traverse_sitemodel(lkcuda_engine *e, int slot, int nops);

}  // namespace lkc
#endif
