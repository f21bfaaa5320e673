// EBCOT Tier-1 decoder for sm_100a: one code-block per warp.
//
// Decoding is serial by nature (every MQ decision feeds the next context), so the warp runs the
// decision loop redundantly in all lanes on warp-uniform data (broadcast loads, no divergence) and
// uses the 32 lanes for what is parallel: clearing the block, folding each finished bit-plane into
// the coefficient plane with coalesced 128-byte row accesses, and the final sign / mid-point /
// dequantisation pass (T1Decoder.cpp:59-99).  Significance, sign, visited and refined state are
// 64-bit row masks in shared memory; the stripe being scanned is held in registers, columns with
// nothing to code are skipped with find-first-set on the masks.
// Reference behaviour: t1.cpp:1337-1448 (block loop), :382-584 (significance), :616-806
// (refinement), :845-1112 (cleanup), mqc.cpp:336-378, 560-595 (MQ decoder).  cblksty == 0 only.
#include "t1_dec_core.cuh"
#include <cstdlib>

namespace j2k {
namespace {

using namespace t1dec;
constexpr int DEC_WARPS = 4;

template <int MINB, bool GEN>
__global__ void __launch_bounds__(32 * DEC_WARPS, MINB)
k_t1_decode(const T1DecBlock* __restrict__ blocks, int nblocks, const T1Tables* __restrict__ tab,
            const uint8_t* __restrict__ seg_arena, const T1DecSeg* __restrict__ segtab, int raw_output) {
    __shared__ DRows s_rows[DEC_WARPS];
    __shared__ ColBuf s_cols[DEC_WARPS];
    __shared__ uint32_t s_state[DEC_WARPS][32];
    __shared__ MqIo s_io[DEC_WARPS];
    __shared__ uint8_t s_zc[4 * 512];
    __shared__ uint8_t s_sc[256];
    __shared__ uint32_t s_mq[96];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) s_zc[i] = tab->zc9[i];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_sc[i] = sc_entry(tab->sc, i);
    if (threadIdx.x < 94) s_mq[threadIdx.x] = mq_state_word(tab->mq, threadIdx.x);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int bi = blockIdx.x * DEC_WARPS + warp;
    if (bi >= nblocks) return;
    const T1DecBlock blk = blocks[bi];
    if (blk.w == 0 || blk.h == 0) return;
    if (blk.npasses == 0 || blk.numbps == 0) {               // block absent from the codestream: all zero
        for (int y = 0; y < blk.h; ++y)
            for (int x = lane; x < blk.w; x += 32) blk.coeff[(size_t)y * blk.pitch + x] = 0;
        return;
    }
    DRows& R = s_rows[warp];
    {
        u64* z = reinterpret_cast<u64*>(&R);
        for (int i = lane; i < (int)(sizeof(DRows) / sizeof(u64)); i += 32) z[i] = 0;
    }
    uint32_t* st = s_state[warp];
    if (lane < T1_NCTX) st[lane] = s_mq[lane == CTX_UNI ? 46 * 2 : (lane == CTX_AGG ? 3 * 2 : (lane == CTX_ZC ? 4 * 2 : 0))];
    __syncwarp();
    decode_block<GEN>(blk, R, s_cols[warp], &s_io[warp], st, s_zc, s_sc, s_mq, seg_arena, segtab, raw_output, lane, 32);
}

}  // namespace

void launch_t1_decode(const T1DecBlock* d_blocks, int nblocks, const uint8_t* seg_arena, const T1DecSeg* d_segs, int generic,
                      int raw_output, cudaStream_t st) {
    if (nblocks <= 0) return;
    // register budget / occupancy trade-off, tunable for experiments: 4 -> 111 regs (16 warps/SM), 5 -> 96 (20), 6 -> 80 (24, the
    // default: best with several frames in flight), 7 -> 72 (28), 8 -> 64 (32)
    static const int variant = [] { const char* e = getenv("J2KCUDA_DEC_MINB"); return e ? atoi(e) : 6; }();
    const dim3 grid((nblocks + DEC_WARPS - 1) / DEC_WARPS), block(32 * DEC_WARPS);
    const T1Tables* tab = t1_tables_device();
    if (generic) { k_t1_decode<4, true><<<grid, block, 0, st>>>(d_blocks, nblocks, tab, seg_arena, d_segs, raw_output); return; }
    if (variant <= 4) k_t1_decode<4, false><<<grid, block, 0, st>>>(d_blocks, nblocks, tab, seg_arena, d_segs, raw_output);
    else if (variant == 5) k_t1_decode<5, false><<<grid, block, 0, st>>>(d_blocks, nblocks, tab, seg_arena, d_segs, raw_output);
    else if (variant == 6) k_t1_decode<6, false><<<grid, block, 0, st>>>(d_blocks, nblocks, tab, seg_arena, d_segs, raw_output);
    else if (variant == 7) k_t1_decode<7, false><<<grid, block, 0, st>>>(d_blocks, nblocks, tab, seg_arena, d_segs, raw_output);
    else k_t1_decode<8, false><<<grid, block, 0, st>>>(d_blocks, nblocks, tab, seg_arena, d_segs, raw_output);
}

}  // namespace j2k
