// ba_window.cu — the window fast path: sliding-window / batched-window bundle adjustment
// (BASELINE configs 1, 2 and 5) where every problem has at most 32 keyframes.
//
// Same math as ba_kernels.cu (reference citations there and in ba_math.cuh), different data
// movement.  A problem's points are cut into TILES of consecutive points (<= kWinPtCap points,
// <= kWinObsCap observations).  One CTA owns one tile and keeps every per-observation
// quantity of the tile in shared memory, so that per LM trial HBM sees only
//     linearize_schur : observations (14 B) + points (24 B) + the tile's pair lists (4 B per pair, built once at upload by
//                       ba_setup.cu from the per-point visibility masks) in, per-point 72 B + per-tile partial
//                       reduced-system blocks out
//     update          : observations + points + per-point 72 B in, points out
// and no per-observation record and no materialised 6x3 cross block exists in global memory.
//
// Staging: the contiguous inputs of a tile (points, per-point records, the problem's R|t tables and pose steps) arrive by
// TMA bulk copies (cp.async.bulk, completion on an mbarrier) issued by one thread, in the layout they have in global
// memory; the observation words are coalesced 4 / 8 / 2-byte loads into registers.
//
// Shared-memory layout of the records: structure of arrays (six 16-byte chunks per
// observation), observations ordered KEYFRAME-major inside the tile.  The pair phase of block
// (I, J) then reads nearly consecutive slots of keyframe I and of keyframe J, which keeps the
// 16-byte gathers of a quarter warp on distinct banks (point-major order made them collide
// 2.3-fold, and the pair phase is bound by shared-memory wavefronts, not by the FP64 pipe).
//
// Scatter-free and deterministic: a sub-warp of 8 lanes owns one 6x6 block of S at a time,
// every lane keeps 36 + 6 fp64 accumulators in registers over "its" observation pairs, the 8
// lanes are folded with a fixed butterfly, and the per-tile partial blocks are summed in tile
// order by k_win_finalize.  No atomics on floating-point data anywhere.
#include "ba_math.cuh"

namespace visoba {

namespace {

constexpr int kSub = 8;                         // lanes per S block
constexpr int kSubsPerCta = kWinThreads / kSub; // blocks in flight per CTA
constexpr int kWarps = kWinThreads / 32;
constexpr int kObsPerThread = (kWinObsCap + kWinThreads - 1) / kWinThreads;
// points and per-point records live in shared memory in the layout of global memory, [point][3] and [point][9], so that one
// bulk copy (TMA, cp.async.bulk + mbarrier) stages a tile's slice; a tile may start at an odd point, i.e. 8 bytes off the
// 16-byte alignment the copy needs: it then starts one double early and the accessors skip that lead slot.  Strides 3 and 9
// doubles are odd multiples of a bank pair: a half-warp of consecutive points hits 16 different bank pairs.
constexpr int kPtDoubles = 3 * kWinPtCap + 2;   // + lead slot + padding to a multiple of 16 bytes
constexpr int kInfoDoubles = 9 * kWinPtCap + 2;

__host__ __device__ constexpr size_t align16(size_t x) { return (x + 15) & ~(size_t)15; }

// ---- shared memory of k_win_linearize_schur (byte offsets) ----
constexpr size_t kLsRec = 0;                                                   // double2 [6][kWinObsCap]
constexpr size_t kLsPt = kLsRec + 6 * sizeof(double2) * kWinObsCap;            // double [lead + np][3]
constexpr size_t kLsInfo = kLsPt + align16(sizeof(double) * kPtDoubles);       // double [np][9]
constexpr size_t kLsRt = kLsInfo + align16(sizeof(double) * kInfoDoubles);     // double [kWinMaxCams][12]
constexpr int kMaxBlkPtr = (kSmallMaxN / 6) * (kSmallMaxN / 6 + 1) / 2 + 1;
constexpr size_t kLsBlkPtr = kLsRt + sizeof(double) * 12 * kWinMaxCams;        // u32 [kMaxBlkPtr]
constexpr size_t kLsObs0 = kLsBlkPtr + align16(sizeof(unsigned) * kMaxBlkPtr); // u16 [kWinPtCap + 1]
constexpr size_t kLsSlot = kLsObs0 + align16(sizeof(unsigned short) * (kWinPtCap + 1)); // u16 [kWinObsCap]
constexpr size_t kLsRed = kLsSlot + align16(sizeof(unsigned short) * kWinObsCap);       // double [kWarps]
constexpr size_t kLsBar = align16(kLsRed + sizeof(double) * kWarps);                    // mbarrier of the bulk copies
constexpr size_t kLsBytes = kLsBar + 16;

// ---- shared memory of k_win_update ----
constexpr size_t kUsTmp = 0;                                                   // double2 [2][kWinObsCap]
constexpr size_t kUsPt = kUsTmp + 2 * sizeof(double2) * kWinObsCap;            // double [lead + np][3]
constexpr size_t kUsInfo = kUsPt + align16(sizeof(double) * kPtDoubles);       // double [lead + np][9]
constexpr size_t kUsRt = kUsInfo + align16(sizeof(double) * kInfoDoubles);     // double [kWinMaxCams][12] current
constexpr size_t kUsRt2 = kUsRt + sizeof(double) * 12 * kWinMaxCams;           // candidate
constexpr size_t kUsDp = kUsRt2 + sizeof(double) * 12 * kWinMaxCams;           // double [kWinMaxCams][6]
constexpr size_t kUsObs0 = kUsDp + sizeof(double) * 6 * kWinMaxCams;           // u16 [kWinPtCap + 1]
constexpr size_t kUsSlot = kUsObs0 + align16(sizeof(unsigned short) * (kWinPtCap + 1)); // u16 [kWinObsCap]
constexpr size_t kUsRed = kUsSlot + align16(sizeof(unsigned short) * kWinObsCap);       // double [2][kWarps]
constexpr size_t kUsBar = align16(kUsRed + sizeof(double) * 2 * kWarps);
constexpr size_t kUsBytes = kUsBar + 16;

// ---- TMA bulk copies global -> shared, completion on an mbarrier (one per CTA, one phase: a CTA stages one tile) ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); // the async proxy must see the initialised barrier
}
__device__ __forceinline__ void mbar_expect_tx(void* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, void* bar)
{ // dst, src 16-byte aligned, bytes a multiple of 16
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src),
                 "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, unsigned parity)
{
    unsigned done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!done);
}
__device__ __forceinline__ unsigned round16(unsigned x) { return (x + 15u) & ~15u; }

// fixed-shape sum over the CTA: shuffles inside a warp, then the warp totals in warp order
__device__ __forceinline__ double cta_sum(double v, double* red, int tid)
{
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if ((tid & 31) == 0) red[tid >> 5] = v;
    __syncthreads();
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) s += red[w];
    return s;
}

// the tile of this CTA: blockIdx.x, or its entry in the compacted list of a straggler trial
__device__ __forceinline__ int tile_index(const DevProblem& P)
{
    return P.tile_map ? __ldg(P.tile_map + blockIdx.x) : (int)blockIdx.x;
}

__device__ __forceinline__ WinTile load_tile(const DevProblem& P, int tile)
{
    const int4* src = reinterpret_cast<const int4*>(P.tiles + tile);
    const int4 a = __ldg(src), b = __ldg(src + 1), c = __ldg(src + 2), d = __ldg(src + 3);
    WinTile t;
    t.prob = a.x, t.p0 = a.y, t.np = a.z, t.o0 = a.w;
    t.no = b.x, t.cam0 = b.y, t.ncam = b.z, t.nf = b.w;
    t.part_off = ((long long)(unsigned)c.x) | ((long long)c.y << 32);
    t.blkptr_off = ((long long)(unsigned)c.z) | ((long long)c.w << 32);
    t.list_off = ((long long)(unsigned)d.x) | ((long long)d.y << 32);
    t.pad = 0;
    return t;
}

// global -> shared without a register round trip (LDGSTS): a tile's inputs are many short strided copies, and a
// load-then-store loop leaves one miss in flight per thread; these all go out back to back and are awaited once
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

} // namespace

// ------------------------------------------------------------------------------------------
// W1 linearize + Schur, one CTA per tile
//   A1 (thread per observation)  projection, residual, Huber weight, 2x3 point Jacobian
//   A2 (thread per point)        H_ll, b_l, 3x3 Cholesky of H_ll + lambda I  -> L^-1, L^-1 b_l
//   A3 (thread per observation)  G~ = F w Jl L^-T,  v~ = F (w r + G zb)
//   B  (8 lanes per S block)     S_IJ -= Jp_a^T G_a G_b^T Jp_b, S_II += w Jp^T Jp, bs_I -= Jp^T v~
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWinThreads, kWinLsCtas) k_win_linearize_schur(DevProblem P, int init_pass, double lambda_override)
{
    extern __shared__ __align__(16) unsigned char sm[];
    double2* const rec = reinterpret_cast<double2*>(sm + kLsRec);
    double* const s_pt = reinterpret_cast<double*>(sm + kLsPt);
    double* const s_info = reinterpret_cast<double*>(sm + kLsInfo);
    double* const s_rt = reinterpret_cast<double*>(sm + kLsRt);
    unsigned int* const s_bp = reinterpret_cast<unsigned int*>(sm + kLsBlkPtr);
    unsigned short* const s_obs0 = reinterpret_cast<unsigned short*>(sm + kLsObs0);
    unsigned short* const s_slot = reinterpret_cast<unsigned short*>(sm + kLsSlot);
    double* const s_red = reinterpret_cast<double*>(sm + kLsRed);
#define REC(c, o) rec[(c) * kWinObsCap + (o)]

    const int tid = threadIdx.x;
    const int tile = tile_index(P);
    const WinTile t = load_tile(P, tile);
    void* const bar = sm + kLsBar;
    if (tid == 0) mbar_init(bar, 1);
    // the tile's points start 8 bytes off a 16-byte boundary when its first point is odd: the bulk copy starts one double early
    const int lead_pt = t.p0 & 1;
    const double* const pt_s = s_pt + lead_pt; // point p, coordinate k: pt_s[3 p + k]

    // ---- A0: everything the tile needs from HBM, issued at once, and what does not depend on the LM state
    // before the state is read.  Thread <-> slot: the observations of the tile in keyframe-major order, so that
    // consecutive threads share the keyframe (R|t reads broadcast) and touch consecutive record slots ----
    unsigned int meta[kObsPerThread];
    float2 uv[kObsPerThread];
    unsigned short slot_r[kObsPerThread];
#pragma unroll
    for (int i = 0; i < kObsPerThread; ++i) {
        const int o = tid + i * kWinThreads;
        meta[i] = 0u, uv[i] = make_float2(0.f, 0.f), slot_r[i] = 0;
        if (o < t.no) {
            meta[i] = __ldg(P.win_meta + t.o0 + o);
            uv[i] = __ldg(P.win_uv + t.o0 + o);
            slot_r[i] = __ldg(P.win_slot + t.o0 + o);
        }
    }
    const int nblk_all = t.nf * (t.nf + 1) / 2;
    { // the tile's pair lists are read once, in phase B, a few rounds ahead of their use: a round is shorter than an
      // HBM miss (ncu: 12 % of the kernel's stall samples sat on that load), so start the lines towards L2 now
        const unsigned int* lp = P.win_pairs + t.list_off;
        const int words = (int)__ldg(P.win_blkptr + t.blkptr_off + (init_pass ? t.nf : nblk_all));
        for (int w = tid * 32; w < words; w += kWinThreads * 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(lp + w));
    }
    for (int k = tid; k <= nblk_all; k += kWinThreads) cp_async4(s_bp + k, P.win_blkptr + t.blkptr_off + k);
    const int obs0_r = tid <= t.np ? __ldg(P.pt_ptr + t.p0 + tid) - t.o0 : 0; // kWinPtCap < kWinThreads
    const LmState& st = P.lm[t.prob];
    if (!st.active && lambda_override < 0.0) {
        cp_async_wait_all();
        return;
    }
    const int cur = st.cur;
    const double lambda = init_pass ? 0.0 : (lambda_override >= 0.0 ? lambda_override : st.lambda);
    {
        if (tid == 0) { // TMA: the tile's points and its problem's R|t, two bulk copies, one mbarrier
            const double* gp = P.pts[cur] + 3 * (size_t)t.p0 - lead_pt;
            const double* grt = P.cam_rt[cur] + 12 * (size_t)t.cam0; // 96 bytes per pose: always aligned
            const unsigned b_pt = round16(8u * (unsigned)(3 * t.np + lead_pt)), b_rt = 96u * (unsigned)t.ncam;
            mbar_expect_tx(bar, b_pt + b_rt);
            bulk_g2s(s_pt, gp, b_pt, bar);
            bulk_g2s(s_rt, grt, b_rt, bar);
        }
#pragma unroll
        for (int i = 0; i < kObsPerThread; ++i) {
            const int o = tid + i * kWinThreads;
            if (o < t.no) s_slot[o] = slot_r[i];
        }
        if (tid <= t.np) s_obs0[tid] = (unsigned short)obs0_r;
    }
    cp_async_wait_all();
    __syncthreads();
    mbar_wait(bar, 0);

    // ---- A1 ----
    double cost = 0.0;
#pragma unroll
    for (int i = 0; i < kObsPerThread; ++i) {
        const int o = tid + i * kWinThreads;
        if (o < t.no) {
            const int c = meta[i] & 0xff, p = (meta[i] >> 8) & 0xff, sl = o;
            ObsGeom g;
            obs_geometry(P, s_rt + 12 * c, pt_s[3 * p], pt_s[3 * p + 1], pt_s[3 * p + 2], uv[i], g);
            cost += g.rho;
            REC(0, sl) = make_double2(g.xn, g.yn);
            REC(1, sl) = make_double2(g.iz, g.w);
            REC(2, sl) = make_double2(g.jl0[0], g.jl0[1]);
            REC(3, sl) = make_double2(g.jl0[2], g.jl1[0]);
            REC(4, sl) = make_double2(g.jl1[1], g.jl1[2]);
            REC(5, sl) = make_double2(g.ru, g.rv);
        }
    }
    if (init_pass) { // cost of the start state: fixed-tree sum of the tile
        const double c = cta_sum(cost, s_red, tid);
        if (tid == 0) {
            P.tile_sums[2 * (size_t)tile] = c;
            P.tile_sums[2 * (size_t)tile + 1] = 0.0;
        }
    }
    __syncthreads();

    // ---- A2 ----
    double maxdiag = 0.0; // init pass: this thread's share of max diag(H_ll)
    for (int p = tid; p < t.np; p += kWinThreads) {
        double h00 = 0, h10 = 0, h11 = 0, h20 = 0, h21 = 0, h22 = 0, b0 = 0, b1 = 0, b2 = 0;
        for (int o = s_obs0[p]; o < s_obs0[p + 1]; ++o) {
            const int sl = s_slot[o];
            const double w = REC(1, sl).y;
            const double2 j01 = REC(2, sl), j2a = REC(3, sl), jb = REC(4, sl), r = REC(5, sl);
            const double a0 = j01.x, a1 = j01.y, a2 = j2a.x, c0 = j2a.y, c1 = jb.x, c2 = jb.y;
            h00 += w * (a0 * a0 + c0 * c0);
            h10 += w * (a1 * a0 + c1 * c0);
            h11 += w * (a1 * a1 + c1 * c1);
            h20 += w * (a2 * a0 + c2 * c0);
            h21 += w * (a2 * a1 + c2 * c1);
            h22 += w * (a2 * a2 + c2 * c2);
            b0 -= w * (a0 * r.x + c0 * r.y);
            b1 -= w * (a1 * r.x + c1 * r.y);
            b2 -= w * (a2 * r.x + c2 * r.y);
        }
        double i00 = 0, i10 = 0, i11 = 0, i20 = 0, i21 = 0, i22 = 0, z0 = 0, z1 = 0, z2 = 0;
        if (init_pass) { // g2o computeLambdaInit: point part of max diag(H); max is order-free
            maxdiag = fmax(maxdiag, fmax(h00, fmax(h11, h22)));
        } else {
            h00 += lambda, h11 += lambda, h22 += lambda;
            const double l00 = sqrt(h00);
            i00 = 1.0 / l00;
            const double l10 = h10 * i00, l20 = h20 * i00;
            const double l11 = sqrt(h11 - l10 * l10);
            i11 = 1.0 / l11;
            const double l21 = (h21 - l20 * l10) * i11;
            const double l22 = sqrt(h22 - l20 * l20 - l21 * l21);
            i22 = 1.0 / l22;
            i10 = -l10 * i00 * i11;
            i21 = -l21 * i11 * i22;
            i20 = -(l20 * i00 + l21 * i10) * i22;
            z0 = i00 * b0;
            z1 = i10 * b0 + i11 * b1;
            z2 = i20 * b0 + i21 * b1 + i22 * b2;
        }
        const double v[9] = { i00, i10, i11, i20, i21, i22, z0, z1, z2 };
        double* gi = P.ptinfo + kPtInfoDoubles * (size_t)(t.p0 + p);
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            s_info[9 * p + k] = v[k];
            gi[k] = v[k];
        }
    }
    if (init_pass) { // one atomic per warp instead of one per point (2000 per window on one address): max is order-free
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) maxdiag = fmax(maxdiag, __shfl_xor_sync(0xffffffffu, maxdiag, off));
        if ((tid & 31) == 0 && maxdiag > 0.0) atomicMax(P.maxdiag_bits + t.prob, (unsigned long long)__double_as_longlong(maxdiag));
    }
    __syncthreads();

    // ---- A3 ----
#pragma unroll
    for (int i = 0; i < kObsPerThread; ++i) {
        const int o = tid + i * kWinThreads;
        if (o < t.no) {
            const int p = (meta[i] >> 8) & 0xff, sl = o;
            const double* in = s_info + 9 * p;
            const double i00 = in[0], i10 = in[1], i11 = in[2], i20 = in[3], i21 = in[4], i22 = in[5];
            const double z0 = in[6], z1 = in[7], z2 = in[8];
            const double w = REC(1, sl).y;
            const double2 j01 = REC(2, sl), j2a = REC(3, sl), jb = REC(4, sl), r = REC(5, sl);
            const double g00 = w * (j01.x * i00);
            const double g01 = w * (j01.x * i10 + j01.y * i11);
            const double g02 = w * (j01.x * i20 + j01.y * i21 + j2a.x * i22);
            const double g10 = w * (j2a.y * i00);
            const double g11 = w * (j2a.y * i10 + jb.x * i11);
            const double g12 = w * (j2a.y * i20 + jb.x * i21 + jb.y * i22);
            const double v0 = w * r.x + g00 * z0 + g01 * z1 + g02 * z2;
            const double v1 = w * r.y + g10 * z0 + g11 * z1 + g12 * z2;
            REC(2, sl) = make_double2(P.fx * g00, P.fx * g01);
            REC(3, sl) = make_double2(P.fx * g02, P.fy * g10);
            REC(4, sl) = make_double2(P.fy * g11, P.fy * g12);
            REC(5, sl) = make_double2(P.fx * v0, P.fy * v1);
        }
    }
    __syncthreads();

    // ---- B ----
    const int l8 = tid & (kSub - 1);
    const int g = tid / kSub; // sub-warp index inside the CTA
    const int nf = t.nf;
    const int nblk = init_pass ? nf : nblk_all; // init pass: diagonal blocks only (H_pp)
    const double fx2 = P.fx * P.fx, fy2 = P.fy * P.fy;
    double* part = P.tile_part + t.part_off;
    const unsigned int* __restrict__ pairs = P.win_pairs + t.list_off;
    const int npass = (nblk + kSubsPerCta - 1) / kSubsPerCta;
    // slot -> block.  Init pass: the nf diagonal blocks.  Otherwise the order ba_setup.cu left behind the tile's offsets
    // (longest pair list first), fetched one pass ahead of its use.
    const unsigned int* __restrict__ order = P.win_blkptr + t.blkptr_off + nblk_all + 1;
    int blk_next = (!init_pass && g < nblk) ? (int)__ldg(order + g) : 0;
    for (int pass = 0; pass < npass; ++pass) {
        const int slot = pass * kSubsPerCta + g;
        const bool have = slot < nblk;
        int I = 0, J = 0;
        if (init_pass) I = J = have ? slot : 0;
        else {
            const int b = have ? blk_next : 0; // triangular index I (I + 1) / 2 + J
            I = (int)((sqrtf(8.0f * (float)b + 1.0f) - 1.0f) * 0.5f);
            while ((I + 1) * (I + 2) / 2 <= b) ++I;
            while (I * (I + 1) / 2 > b) --I;
            J = b - I * (I + 1) / 2;
            blk_next = slot + kSubsPerCta < nblk ? (int)__ldg(order + slot + kSubsPerCta) : 0;
        }
        const bool diag = I == J;
        const int blk = I * (I + 1) / 2 + J;
        // the block's pair list inside this tile (built once at upload): points both keyframes see
        const int beg = have ? (int)s_bp[blk] : 0;
        const int cnt = have ? (int)s_bp[blk + 1] - beg : 0;
        const unsigned int* __restrict__ list = pairs + beg;
        int maxcnt = cnt;
        maxcnt = max(maxcnt, __shfl_xor_sync(0xffffffffu, maxcnt, 8));
        maxcnt = max(maxcnt, __shfl_xor_sync(0xffffffffu, maxcnt, 16));
        // pair words are prefetched three rounds ahead (a round is far shorter than an L2 miss)
        unsigned int pr_q0 = l8 < cnt ? __ldg(list + l8) : 0u;
        unsigned int pr_q1 = l8 + kSub < cnt ? __ldg(list + l8 + kSub) : 0u;
        unsigned int pr_q2 = l8 + 2 * kSub < cnt ? __ldg(list + l8 + 2 * kSub) : 0u;

        double v[48];
#pragma unroll
        for (int i = 0; i < 48; ++i) v[i] = 0.0;
        for (int i0 = 0; i0 < maxcnt; i0 += kSub) {
            const int i = i0 + l8;
            const unsigned int pr = pr_q0;
            pr_q0 = pr_q1, pr_q1 = pr_q2;
            pr_q2 = i + 3 * kSub < cnt ? __ldg(list + i + 3 * kSub) : 0u;
            if (i < cnt) {
                const int oa = pr & 0xffffu, ob = pr >> 16;
                const double2 a0 = REC(0, oa), a1 = REC(1, oa), a2 = REC(2, oa), a3 = REC(3, oa), a4 = REC(4, oa);
                const double2 b0 = REC(0, ob), b1 = REC(1, ob), b2 = REC(2, ob), b3 = REC(3, ob), b4 = REC(4, ob);
                // M = -G~_a G~_b^T  (+ w F^2 on the diagonal block);  G~ = (c2.x c2.y c3.x ; c3.y c4.x c4.y)
                double m00 = -fma(a2.x, b2.x, fma(a2.y, b2.y, a3.x * b3.x));
                double m01 = -fma(a2.x, b3.y, fma(a2.y, b4.x, a3.x * b4.y));
                double m10 = -fma(a3.y, b2.x, fma(a4.x, b2.y, a4.y * b3.x));
                double m11 = -fma(a3.y, b3.y, fma(a4.x, b4.x, a4.y * b4.y));
                if (diag) {
                    m00 = fma(a1.y, fx2, m00);
                    m11 = fma(a1.y, fy2, m11);
                }
                // J~ rows (viso.cpp:653-655 without fx, fy):
                //   r0 = [iz, 0, -x iz, -x y, 1 + x^2, -y]    r1 = [0, iz, -y iz, -(1 + y^2), x y, x]
                const double xb = b0.x, yb = b0.y, zb = b1.x;
                const double bxz = xb * zb, byz = yb * zb, bxy = xb * yb, bxx = fma(xb, xb, 1.0), byy = fma(yb, yb, 1.0);
                double t0[6], t1[6];
                t0[0] = m00 * zb, t1[0] = m10 * zb;
                t0[1] = m01 * zb, t1[1] = m11 * zb;
                t0[2] = -fma(m00, bxz, m01 * byz), t1[2] = -fma(m10, bxz, m11 * byz);
                t0[3] = -fma(m00, bxy, m01 * byy), t1[3] = -fma(m10, bxy, m11 * byy);
                t0[4] = fma(m00, bxx, m01 * bxy), t1[4] = fma(m10, bxx, m11 * bxy);
                t0[5] = fma(m01, xb, -(m00 * yb)), t1[5] = fma(m11, xb, -(m10 * yb));
                const double xa = a0.x, ya = a0.y, za = a1.x;
                const double axz = -(xa * za), ayz = -(ya * za), axy = xa * ya, axx = fma(xa, xa, 1.0), ayy = -fma(ya, ya, 1.0);
                const double naxy = -axy, nya = -ya;
#pragma unroll
                for (int c = 0; c < 6; ++c) {
                    v[c] = fma(za, t0[c], v[c]);
                    v[6 + c] = fma(za, t1[c], v[6 + c]);
                    v[12 + c] = fma(axz, t0[c], fma(ayz, t1[c], v[12 + c]));
                    v[18 + c] = fma(naxy, t0[c], fma(ayy, t1[c], v[18 + c]));
                    v[24 + c] = fma(axx, t0[c], fma(axy, t1[c], v[24 + c]));
                    v[30 + c] = fma(nya, t0[c], fma(xa, t1[c], v[30 + c]));
                }
                if (diag) { // rhs: Jp^T v~
                    const double2 a5 = REC(5, oa);
                    v[36] = fma(za, a5.x, v[36]);
                    v[37] = fma(za, a5.y, v[37]);
                    v[38] = fma(axz, a5.x, fma(ayz, a5.y, v[38]));
                    v[39] = fma(naxy, a5.x, fma(ayy, a5.y, v[39]));
                    v[40] = fma(axx, a5.x, fma(axy, a5.y, v[40]));
                    v[41] = fma(nya, a5.x, fma(xa, a5.y, v[41]));
                }
            }
        }
        // fold the 8 lanes of the sub-warp: afterwards lane l8 holds entries 6 l8 .. 6 l8 + 5
#pragma unroll
        for (int off = 4, n = 24; off >= 1; off >>= 1, n >>= 1) {
            const bool upper = (l8 & off) != 0;
#pragma unroll
            for (int i = 0; i < n; ++i) {
                const double send = upper ? v[i] : v[i + n];
                const double keep = upper ? v[i + n] : v[i];
                v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
            }
        }
        if (have && l8 < 7) {
            double* out = part + (size_t)blk * kItemPart + 6 * l8;
#pragma unroll
            for (int i = 0; i < 6; ++i) out[i] = v[i];
        }
    }
#undef REC
}

// ------------------------------------------------------------------------------------------
// W2 finalize: one CTA per problem sums the per-tile partial blocks in tile order, adds lambda on
// the diagonal (g2o setLambda) and writes S_blk / bs (g2o solves H x = b, b = -J^T r).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_win_finalize(DevProblem P, int init_pass, double lambda_override)
{
    const int b = blockIdx.x;
    const LmState& st = P.lm[b];
    if (!st.active && lambda_override < 0.0) return;
    const double lambda = init_pass ? 0.0 : (lambda_override >= 0.0 ? lambda_override : st.lambda);
    const int nf = P.prob_nfree[b];
    const int nblk = nf * (nf + 1) / 2;
    const int t0 = P.prob_tile_ptr[b], t1 = P.prob_tile_ptr[b + 1];
    const double* __restrict__ part = P.tile_part + P.prob_part_off[b];
    const size_t stride = (size_t)nblk * kItemPart;
    const int tri0 = P.prob_tri0[b];
    // gridDim.y chunks per problem: a lone window's partials would otherwise be summed by ONE CTA (config 2: 8 MB)
    for (int idx = blockIdx.y * blockDim.x + threadIdx.x; idx < nblk * kItemPart; idx += gridDim.y * blockDim.x) {
        const int blk = idx / kItemPart, e = idx - blk * kItemPart;
        const int cam = P.block_cam[tri0 + blk];
        if (cam < 0 && (e >= 36 || init_pass)) continue;
        double s = 0.0;
        for (int t = 0; t < t1 - t0; ++t) s += part[(size_t)t * stride + idx];
        if (e < 36) {
            if (cam >= 0 && (e % 7) == 0) s += lambda;
            P.S_blk[36 * (size_t)(tri0 + blk) + e] = s;
        } else {
            P.bs[6 * (size_t)cam + (e - 36)] = -s;
        }
    }
}

// ------------------------------------------------------------------------------------------
// W3 update, one CTA per tile: back substitution, candidate points and their robust cost.
//   U1 (thread per observation)  everything of the observation is RECOMPUTED from the current
//                                state (cheaper than 96 B/obs of HBM traffic):  t = Jp d_p,
//                                its share of  sum_a G_a^T Jp_a d_p  and of g2o's computeScale
//   U2 (thread per point)        d_l = L^-T (zb - sum),  P' = P + d_l
//   U3 (thread per observation)  robust cost at the candidate state
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWinThreads, 5) k_win_update(DevProblem P)
{
    extern __shared__ __align__(16) unsigned char sm[];
    double2* const tmp = reinterpret_cast<double2*>(sm + kUsTmp);
    double* const s_pt = reinterpret_cast<double*>(sm + kUsPt);
    double* const s_info = reinterpret_cast<double*>(sm + kUsInfo);
    double* const s_rt = reinterpret_cast<double*>(sm + kUsRt);
    double* const s_rt2 = reinterpret_cast<double*>(sm + kUsRt2);
    double* const s_dp = reinterpret_cast<double*>(sm + kUsDp);
    unsigned short* const s_obs0 = reinterpret_cast<unsigned short*>(sm + kUsObs0);
    unsigned short* const s_slot = reinterpret_cast<unsigned short*>(sm + kUsSlot);
    double* const s_red = reinterpret_cast<double*>(sm + kUsRed);

    const int tid = threadIdx.x;
    const int tile = tile_index(P);
    const WinTile t = load_tile(P, tile);
    // everything that does not depend on the LM state goes out before the state is read (one dependent miss less)
    unsigned int meta[kObsPerThread];
    float2 uv[kObsPerThread];
    unsigned short slot_r[kObsPerThread];
#pragma unroll
    for (int i = 0; i < kObsPerThread; ++i) {
        const int o = tid + i * kWinThreads;
        meta[i] = 0u, uv[i] = make_float2(0.f, 0.f), slot_r[i] = 0;
        if (o < t.no) {
            meta[i] = __ldg(P.win_meta + t.o0 + o);
            uv[i] = __ldg(P.win_uv + t.o0 + o);
            slot_r[i] = __ldg(P.win_slot + t.o0 + o);
        }
    }
    // TMA staging (cp.async.bulk + mbarrier).  Barrier 0: what does not depend on the LM state (per-point records, pose
    // steps), issued before the state is read; barrier 1: points and both R|t tables of the problem's current / candidate
    // buffers.  A tile that starts at an odd point starts its copies one point-record early (16-byte alignment).
    void* const bar0 = sm + kUsBar;
    void* const bar1 = sm + kUsBar + 8;
    const int lead = t.p0 & 1; // in points: 72 and 24 bytes per point are 8 bytes off the alignment for an odd point
    double* const pt_s = s_pt + lead;           // point p, coordinate k: pt_s[3 p + k]   (lead: one double)
    const double* const info_s = s_info + lead; // point p, entry k:      info_s[9 p + k]
    if (tid == 0) {
        mbar_init(bar0, 1);
        mbar_init(bar1, 1);
        const double* gi = P.ptinfo + kPtInfoDoubles * (size_t)t.p0 - lead;
        const double* gd = P.dp + 6 * (size_t)t.cam0; // 48 bytes per pose: always aligned
        const unsigned b_i = round16(8u * (unsigned)(9 * t.np + lead)), b_d = 48u * (unsigned)t.ncam;
        mbar_expect_tx(bar0, b_i + b_d);
        bulk_g2s(s_info, gi, b_i, bar0);
        bulk_g2s(s_dp, gd, b_d, bar0);
    }
    const int obs0_r = tid <= t.np ? __ldg(P.pt_ptr + t.p0 + tid) - t.o0 : 0; // kWinPtCap < kWinThreads
    const LmState& st = P.lm[t.prob];
    if (!st.active) { // no CTA may exit with a copy into its shared memory in flight
        __syncthreads();
        mbar_wait(bar0, 0);
        return;
    }
    const int cur = st.cur, cand = cur ^ 1;
    const double lambda = st.lambda;
    {
        if (tid == 0) {
            const double* gp = P.pts[cur] + 3 * (size_t)t.p0 - lead;
            const double* g1 = P.cam_rt[cur] + 12 * (size_t)t.cam0;
            const double* g2 = P.cam_rt[cand] + 12 * (size_t)t.cam0;
            const unsigned b_p = round16(8u * (unsigned)(3 * t.np + lead)), b_r = 96u * (unsigned)t.ncam;
            mbar_expect_tx(bar1, b_p + 2 * b_r);
            bulk_g2s(s_pt, gp, b_p, bar1);
            bulk_g2s(s_rt, g1, b_r, bar1);
            bulk_g2s(s_rt2, g2, b_r, bar1);
        }
#pragma unroll
        for (int i = 0; i < kObsPerThread; ++i) {
            const int o = tid + i * kWinThreads;
            if (o < t.no) s_slot[o] = slot_r[i];
        }
        if (tid <= t.np) s_obs0[tid] = (unsigned short)obs0_r;
    }
    __syncthreads();
    mbar_wait(bar0, 0);
    mbar_wait(bar1, 0);

    // ---- U1 ----
#pragma unroll
    for (int i = 0; i < kObsPerThread; ++i) {
        const int o = tid + i * kWinThreads;
        if (o < t.no) {
            const int c = meta[i] & 0xff, p = (meta[i] >> 8) & 0xff;
            ObsGeom g;
            obs_geometry(P, s_rt + 12 * c, pt_s[3 * p], pt_s[3 * p + 1], pt_s[3 * p + 2], uv[i], g);
            const double* in = info_s + 9 * p;
            const double i00 = in[0], i10 = in[1], i11 = in[2], i20 = in[3], i21 = in[4], i22 = in[5];
            const double g00 = P.fx * (g.w * (g.jl0[0] * i00));
            const double g01 = P.fx * (g.w * (g.jl0[0] * i10 + g.jl0[1] * i11));
            const double g02 = P.fx * (g.w * (g.jl0[0] * i20 + g.jl0[1] * i21 + g.jl0[2] * i22));
            const double g10 = P.fy * (g.w * (g.jl1[0] * i00));
            const double g11 = P.fy * (g.w * (g.jl1[0] * i10 + g.jl1[1] * i11));
            const double g12 = P.fy * (g.w * (g.jl1[0] * i20 + g.jl1[1] * i21 + g.jl1[2] * i22));
            double ja0[6], ja1[6];
            jtilde(g.xn, g.yn, g.iz, ja0, ja1);
            const double* d = s_dp + 6 * c;
            double t0 = 0.0, t1 = 0.0;
#pragma unroll
            for (int k = 0; k < 6; ++k) {
                t0 += ja0[k] * d[k];
                t1 += ja1[k] * d[k];
            }
            // pose part of computeScale: d_p . b_p = -sum t~ . (F w r)
            const double sc = -(t0 * (P.fx * (g.w * g.ru)) + t1 * (P.fy * (g.w * g.rv)));
            tmp[o] = make_double2(g00 * t0 + g10 * t1, g01 * t0 + g11 * t1);
            tmp[kWinObsCap + o] = make_double2(g02 * t0 + g12 * t1, sc);
        }
    }
    __syncthreads();

    // ---- U2 ----
    double scale = 0.0;
    for (int p = tid; p < t.np; p += kWinThreads) {
        const double* in = info_s + 9 * p;
        const double i00 = in[0], i10 = in[1], i11 = in[2], i20 = in[3], i21 = in[4], i22 = in[5];
        const double z0 = in[6], z1 = in[7], z2 = in[8];
        double y0 = z0, y1 = z1, y2 = z2, sc = 0.0;
        for (int o = s_obs0[p]; o < s_obs0[p + 1]; ++o) {
            const int sl = s_slot[o];
            const double2 q0 = tmp[sl], q1 = tmp[kWinObsCap + sl];
            y0 -= q0.x;
            y1 -= q0.y;
            y2 -= q1.x;
            sc += q1.y;
        }
        const double d0 = i00 * y0 + i10 * y1 + i20 * y2;
        const double d1 = i11 * y1 + i21 * y2;
        const double d2 = i22 * y2;
        sc += y0 * z0 + y1 * z1 + y2 * z2 + lambda * (d0 * d0 + d1 * d1 + d2 * d2);
        scale += sc;
        pt_s[3 * p] += d0, pt_s[3 * p + 1] += d1, pt_s[3 * p + 2] += d2;
    }
    __syncthreads();
    { // candidate points out, coalesced
        double* __restrict__ gp = P.pts[cand] + 3 * (size_t)t.p0;
        for (int i = tid; i < 3 * t.np; i += kWinThreads) gp[i] = pt_s[i];
    }

    // ---- U3 ----
    double cost = 0.0;
#pragma unroll
    for (int i = 0; i < kObsPerThread; ++i) {
        const int o = tid + i * kWinThreads;
        if (o < t.no) {
            const int c = meta[i] & 0xff, p = (meta[i] >> 8) & 0xff;
            const double* rt = s_rt2 + 12 * c;
            const double X = pt_s[3 * p], Y = pt_s[3 * p + 1], Z = pt_s[3 * p + 2];
            const double x = rt[0] * X + rt[1] * Y + rt[2] * Z + rt[9];
            const double y = rt[3] * X + rt[4] * Y + rt[5] * Z + rt[10];
            const double z = rt[6] * X + rt[7] * Y + rt[8] * Z + rt[11];
            const double iz = 1.0 / z;
            const double ru = P.fx * (x * iz) + P.cx - (double)uv[i].x;
            const double rv = P.fy * (y * iz) + P.cy - (double)uv[i].y;
            double w;
            cost += huber_rho(ru * ru + rv * rv, P.huber_delta, w);
        }
    }
    const double c = cta_sum(cost, s_red, tid);
    const double sc = cta_sum(scale, s_red + kWarps, tid);
    if (tid == 0) {
        P.tile_sums[2 * (size_t)tile] = c;
        P.tile_sums[2 * (size_t)tile + 1] = sc;
    }
}

// per-problem totals of the tile sums, in tile order
__global__ void k_win_prob_sums(DevProblem P, int even_if_inactive)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= P.n_problems) return;
    if (!P.lm[b].active && !even_if_inactive) return;
    double s0 = 0.0, s1 = 0.0;
    for (int t = P.prob_tile_ptr[b]; t < P.prob_tile_ptr[b + 1]; ++t) {
        s0 += P.tile_sums[2 * (size_t)t];
        s1 += P.tile_sums[2 * (size_t)t + 1];
    }
    P.prob_sums[2 * (size_t)b] = s0;
    P.prob_sums[2 * (size_t)b + 1] = s1;
}

// ------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------
size_t win_shared_bytes() { return kLsBytes; }

static bool win_configure()
{
    static PerDeviceOnce once;
    return once.run([] {
        cudaError_t e1 = cudaFuncSetAttribute(k_win_linearize_schur, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kLsBytes);
        cudaError_t e2 = cudaFuncSetAttribute(k_win_update, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kUsBytes);
        return e1 == cudaSuccess && e2 == cudaSuccess;
    });
}

int launch_win_linearize_schur(const DevProblem& P, int init_pass, double lambda_override, cudaStream_t st)
{
    if (P.n_tiles <= 0 || !win_configure()) return 0;
    k_win_linearize_schur<<<P.tile_map ? P.n_map_tiles : P.n_tiles, kWinThreads, kLsBytes, st>>>(P, init_pass, lambda_override);
    return 1;
}

int launch_win_finalize(const DevProblem& P, int init_pass, double lambda_override, cudaStream_t st)
{
    // enough CTAs to fill the GPU: a batch of windows has them already (one CTA per problem reads its partials at the HBM
    // roof), a few windows get their entries cut into chunks of 256
    const int per = (int)(((long long)P.n_blocks * kItemPart / (P.n_problems > 0 ? P.n_problems : 1) + 255) / 256);
    const int chunks = P.n_problems >= 296 ? 1 : (per < 1 ? 1 : (per > 64 ? 64 : per));
    k_win_finalize<<<dim3((unsigned)P.n_problems, (unsigned)chunks), 256, 0, st>>>(P, init_pass, lambda_override);
    return 1;
}

int launch_win_update(const DevProblem& P, cudaStream_t st)
{
    if (P.n_tiles <= 0 || !win_configure()) return 0;
    k_win_update<<<P.tile_map ? P.n_map_tiles : P.n_tiles, kWinThreads, kUsBytes, st>>>(P);
    return 1;
}

// the tiles of the problems still active, in tile order, and their count at P.n_active[1]: one CTA, a running
// prefix over the problems.  The host sizes the straggler trials' grids with that count.
__global__ void __launch_bounds__(1024) k_win_active_tiles(DevProblem P, int32_t* map)
{
    __shared__ int warp_tot[32];
    __shared__ int base;
    if (threadIdx.x == 0) base = 0;
    __syncthreads();
    for (int b0 = 0; b0 < P.n_problems; b0 += blockDim.x) {
        const int b = b0 + threadIdx.x;
        int t0 = 0, cnt = 0;
        if (b < P.n_problems && P.lm[b].active) t0 = P.prob_tile_ptr[b], cnt = P.prob_tile_ptr[b + 1] - t0;
        int inc = cnt; // inclusive prefix inside the warp
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, inc, off);
            if ((threadIdx.x & 31) >= off) inc += v;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
        __syncthreads();
        int before = base;
        for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) before += warp_tot[w];
        for (int k = 0; k < cnt; ++k) map[before + inc - cnt + k] = t0 + k;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) base = before + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) P.n_active[1] = base;
}

int launch_win_active_tiles(const DevProblem& P, int32_t* map, cudaStream_t st)
{
    k_win_active_tiles<<<1, 1024, 0, st>>>(P, map);
    return 1;
}

// the per-problem totals and the LM decision in one launch: one WARP per problem — the lanes fetch the sums of 32 tiles at once
// (one L2 round trip instead of one per eight tiles), lane 0's additions stay sequential in tile order
__global__ void __launch_bounds__(128) k_win_sums_decide(DevProblem P)
{
    const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (b >= P.n_problems) return;
    if (!P.lm[b].active) return;
    double s0 = 0.0, s1 = 0.0;
    const int t1 = P.prob_tile_ptr[b + 1];
    for (int t = P.prob_tile_ptr[b]; t < t1; t += 32) {
        const double2 v = t + lane < t1 ? *reinterpret_cast<const double2*>(P.tile_sums + 2 * (size_t)(t + lane)) : make_double2(0.0, 0.0);
        const int n = min(32, t1 - t);
        for (int k = 0; k < n; ++k) {
            s0 += __shfl_sync(0xffffffffu, v.x, k);
            s1 += __shfl_sync(0xffffffffu, v.y, k);
        }
    }
    if (lane) return;
    P.prob_sums[2 * (size_t)b] = s0;
    P.prob_sums[2 * (size_t)b + 1] = s1;
    lm_decide(P, b);
}

int launch_win_sums_decide(const DevProblem& P, cudaStream_t st)
{
    k_win_sums_decide<<<(P.n_problems + 3) / 4, 128, 0, st>>>(P);
    return 1;
}

int launch_win_prob_sums(const DevProblem& P, int even_if_inactive, cudaStream_t st)
{
    k_win_prob_sums<<<(P.n_problems + 127) / 128, 128, 0, st>>>(P, even_if_inactive);
    return 1;
}

static_assert(kWinLsCtas * (kLsBytes + 1024) <= 227 * 1024, "the resident linearize CTAs of one SM must fit in shared memory");
static_assert(kWinPtCap <= 255 && kWinObsCap <= 65535 && kWinMaxCams <= 32, "meta word packing");
static_assert(kWinPtCap < kWinThreads, "one thread per point-pointer entry at tile entry");

} // namespace visoba
