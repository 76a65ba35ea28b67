// fast.cu — K2: per-cell FAST-9/16 with score, per-cell 3x3 non-max suppression and the
// iniThFAST -> minThFAST retry (reference src/ORBextractor.cc:815-896 calling cv::FAST twice
// per 30-px cell; cv::FAST semantics restated in SURVEY.md §8c).
//
// One CTA per cell, one launch for all cells of all levels of all frames.  The CTA stages the
// cell window (cell + 3-px ring radius on each side) in shared memory, computes for every
// pixel of the detection zone the threshold-independent corner strength
//     S = max( max_arc min_{9 contiguous} (v - ring), max_arc min_{9 contiguous} (ring - v) )
// (a pixel is a FAST corner at threshold t iff S > t, and cv::FAST's score is S - 1), keeps
// S where S > t, and applies the strict 3x3 maximum test inside the zone only — exactly what
// running cv::FAST on the cell sub-image does, because its score rows are zero outside the
// sub-image's 3-px inset.  If no corner survives at iniThFAST the same window is re-evaluated
// at minThFAST.  Survivors are emitted in row-major order through a ballot/popc ordered
// compaction, packed x | y<<12 | score<<24 with coordinates relative to (minBorderX,
// minBorderY) as the reference stores them (:889-890).
#include "orbfe_internal.cuh"
#include <cstddef>
#include <cstring>
#include <cuda.h>

namespace orbfe {

namespace {

constexpr int kThreads = 128;
constexpr int kWarpWinMax = 48;          // largest window side the warp-per-cell kernel takes
constexpr int WIN = kFastMaxWin;        // max window side
constexpr int ZONE = WIN - 6;           // max zone side

// cv::FAST ring order, (dx,dy): (0,3) (1,3) (2,2) (3,1) (3,0) (3,-1) (2,-2) (1,-3) (0,-3)
// (-1,-3) (-2,-2) (-3,-1) (-3,0) (-3,1) (-2,2) (-1,3)

// a / b for 0 <= a < 2^16, 0 < b: exact, because (a + 0.5) / b is at least 0.5 / b away from
// an integer while the float error is a few ulp
__device__ __forceinline__ int small_div(int a, float rcp_b) { return (int)(((float)a + 0.5f) * rcp_b); }

// corner strength S of the pixel at p (row stride `st`), or 0 when S <= t is already certain
// QUICK: early exits on the four axis / diagonal pairs (worth it when most pixels fail them; the
// warp kernel's survivors already passed the unsigned form of those tests and skip them)
template <bool QUICK>
__device__ __forceinline__ int corner_strength(const uint8_t* p, int st, int t) {
    const int v = p[0];
    int r[16];
    // opposing pairs first: a 9-arc contains at least one pixel of every pair (k, k+8)
    r[0] = p[3 * st];  r[8] = p[-3 * st];
    r[4] = p[3];  r[12] = p[-3];
    if (QUICK) {
        const int lo = v - t, hi = v + t;
        int m = ((r[0] < lo || r[8] < lo) ? 1 : 0) | ((r[0] > hi || r[8] > hi) ? 2 : 0);
        if (!m) return 0;
        m &= ((r[4] < lo || r[12] < lo) ? 1 : 0) | ((r[4] > hi || r[12] > hi) ? 2 : 0);
        if (!m) return 0;
        r[2] = p[2 * st + 2];  r[10] = p[-2 * st - 2];
        m &= ((r[2] < lo || r[10] < lo) ? 1 : 0) | ((r[2] > hi || r[10] > hi) ? 2 : 0);
        if (!m) return 0;
        r[6] = p[-2 * st + 2];  r[14] = p[2 * st - 2];
        m &= ((r[6] < lo || r[14] < lo) ? 1 : 0) | ((r[6] > hi || r[14] > hi) ? 2 : 0);
        if (!m) return 0;
    } else {
        r[2] = p[2 * st + 2];  r[10] = p[-2 * st - 2];
        r[6] = p[-2 * st + 2];  r[14] = p[2 * st - 2];
    }
    r[1] = p[3 * st + 1];   r[9] = p[-3 * st - 1];
    r[3] = p[st + 3];       r[11] = p[-st - 3];
    r[5] = p[-st + 3];      r[13] = p[st - 3];
    r[7] = p[-3 * st + 1];  r[15] = p[3 * st - 1];
    // (256 + v - r) in the low half, (256 + r - v) in the high half: one IMAD per ring pixel,
    // both halves stay positive so signed 16-bit min/max order them like the differences.
    const uint32_t bias = (uint32_t)(256 + v) | ((uint32_t)(256 - v) << 16);
    uint32_t d[16];
#pragma unroll
    for (int k = 0; k < 16; k++) d[k] = bias + (uint32_t)r[k] * 0xFFFFu;
    // minima over ring[k+1..k+8] for even k (the 8 interior pixels shared by the two 9-arcs
    // starting at k and k+1), by doubling on the odd positions only
    // max(min(m8, a), min(m8, b)) = min(m8, max(a, b)), and the three-input VIMNMX3 folds m8 = min(m4[j], m4[j+2])
    // into the same instruction: 36 min / max operations per pixel
    uint32_t m2[8], m4[8], u[8];
#pragma unroll
    for (int j = 0; j < 8; j++) m2[j] = __vmins2(d[2 * j + 1], d[(2 * j + 2) & 15]);           // [2j+1, 2j+2]
#pragma unroll
    for (int j = 0; j < 8; j++) m4[j] = __vmins2(m2[j], m2[(j + 1) & 7]);                         // [2j+1 .. 2j+4]
#pragma unroll
    for (int j = 0; j < 8; j++)                                                                   // arcs starting at 2j and 2j+1
        u[j] = __vimin3_s16x2(m4[j], m4[(j + 2) & 7], __vmaxs2(d[2 * j], d[(2 * j + 9) & 15]));
    const uint32_t best = __vmaxs2(__vimax3_s16x2(__vimax3_s16x2(u[0], u[1], u[2]), __vimax3_s16x2(u[3], u[4], u[5]), u[6]), u[7]);
    const int A = (int)(best & 0xFFFFu) - 256, B = (int)(best >> 16) - 256;
    const int S = max(A, B);
    return S > t ? S : 0;
}

__global__ void __launch_bounds__(kThreads)
fast_cells_kernel(const __grid_constant__ FrameSetGeom g, const uint8_t* __restrict__ pyr,
                  int* __restrict__ cellc, uint32_t* __restrict__ cellk, int cell0) {
    __shared__ uint8_t win[WIN][WIN + 8];
    __shared__ uint8_t sc[ZONE + 2][ZONE + 2 + 2];
    __shared__ int warp_cnt[kThreads / 32];
    __shared__ int s_total;

    const int cell = blockIdx.x + cell0, frame = blockIdx.y + g.frame0;
    int l = 0;
    while (l + 1 < g.nlevels && cell >= g.lv[l + 1].cellBase) l++;
    const LevelGeom& L = g.lv[l];
    if (L.wCell + 6 <= kWarpWinMax && L.hCell + 6 <= kWarpWinMax) return;   // done by fast_cells_warp_kernel
    const int c = cell - L.cellBase;
    const int ci = c / L.nCols, cj = c - ci * L.nCols;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int* out_cnt = cellc + (long long)frame * g.ncells + cell;
    uint32_t* out = cellk + (long long)frame * g.cellSlots + L.cellSlotOff + (long long)c * L.cellCap;

    // cell window — ORBextractor.cc:846-869 (integer-valued floats there)
    const int iniY = L.minBY + ci * L.hCell, iniX = L.minBX + cj * L.wCell;
    int maxY = iniY + L.hCell + 6, maxX = iniX + L.wCell + 6;
    if (maxY > L.maxBY) maxY = L.maxBY;
    if (maxX > L.maxBX) maxX = L.maxBX;
    const int ww = maxX - iniX, wh = maxY - iniY;
    if (iniY >= L.maxBY - 3 || iniX >= L.maxBX - 6 || ww < 7 || wh < 7) {
        if (tid == 0) *out_cnt = 0;
        return;
    }
    const int zw = ww - 6, zh = wh - 6;
    const uint8_t* src = pyr + L.pyrOff + (long long)frame * L.lvBytes +
                         (long long)iniY * L.pitch + iniX;
    for (int i = tid; i < ww * wh; i += kThreads) {
        const int r = i / ww, cc = i - r * ww;
        win[r][cc] = src[(long long)r * L.pitch + cc];
    }
    for (int i = tid; i < (zh + 2) * (zw + 2); i += kThreads) {   // zero ring of the score map
        const int r = i / (zw + 2), cc = i - r * (zw + 2);
        if (r == 0 || r == zh + 1 || cc == 0 || cc == zw + 1) sc[r][cc] = 0;
    }
    if (tid == 0) s_total = 0;
    __syncthreads();

    const int npx = zw * zh;
    for (int pass = 0; pass < 2; pass++) {
        const int t = pass == 0 ? g.iniTh : g.minTh;
        for (int i = tid; i < npx; i += kThreads) {
            const int r = i / zw, cc = i - r * zw;
            sc[r + 1][cc + 1] = (uint8_t)corner_strength<true>(&win[r + 3][cc + 3], WIN + 8, t);
        }
        __syncthreads();
        // 3x3 strict maximum + ordered (row-major) compaction
        for (int base = 0; base < npx; base += kThreads) {
            const int i = base + tid;
            bool keep = false;
            int r = 0, cc = 0, s = 0;
            if (i < npx) {
                r = i / zw; cc = i - r * zw;
                s = sc[r + 1][cc + 1];
                keep = s > 0 && s > sc[r][cc] && s > sc[r][cc + 1] && s > sc[r][cc + 2] &&
                       s > sc[r + 1][cc] && s > sc[r + 1][cc + 2] && s > sc[r + 2][cc] &&
                       s > sc[r + 2][cc + 1] && s > sc[r + 2][cc + 2];
            }
            const unsigned b = __ballot_sync(0xFFFFFFFFu, keep);
            if (lane == 0) warp_cnt[wid] = __popc(b);
            __syncthreads();
            int pos = s_total;
            for (int w = 0; w < wid; w++) pos += warp_cnt[w];
            pos += __popc(b & ((1u << lane) - 1u));
            if (keep && pos < L.cellCap)
                out[pos] = pack_kp(cc + 3 + cj * L.wCell, r + 3 + ci * L.hCell, s - 1);
            __syncthreads();
            if (tid == 0) {
                int tot = s_total;
                for (int w = 0; w < kThreads / 32; w++) tot += warp_cnt[w];
                s_total = tot;
            }
            __syncthreads();
        }
        if (s_total > 0) break;
    }
    if (tid == 0) *out_cnt = min(s_total, L.cellCap);
}


// ---------------------------------------------------------------------------------------
// Warp-per-cell kernel (all cells whose window fits 48x48, i.e. every level of the usual
// camera sizes).  Stage 1 rejects pixels four at a time with byte-SIMD arithmetic: a 9-arc
// always contains one pixel of each opposing ring pair, so a corner needs, for every pair,
// |ring - v| > t on at least one side; VABSDIFF4 plus a carry-free SWAR compare evaluate that
// for the pairs (0,8), (4,12), (2,10), (6,14) on whole 32-bit words of the aligned window.
// Stage 2 runs the exact polarity-aware test and the corner strength only on the survivors
// (a dense work list in shared memory).  Stage 3 applies the per-cell 3x3 strict maximum and
// marks survivors in a bitmap; popc prefix sums over the bitmap give every survivor its
// row-major output rank, so the emitted order equals cv::FAST's.
// Two layouts: windows up to 40x40 (every cell of the usual camera sizes: cells are 30..34 px
// wide) in 6 KB per warp, which lets 8 CTAs = 32 warps live on an SM, and windows up to 48x48.
constexpr int kWarpsPerCta = 4;

template <int WWIN_> struct S_sc_t { uint8_t b[WWIN_ - 4][WWIN_]; };
// TMA descriptors of the levels (box 64 x window rows), passed by value
// m: box 64 x window rows; m48: box 48 x window rows — a 48-byte box row is exactly the 12-word pitch the stages read,
// so a cell whose window fits it ((iniX & 15) + width <= 48) lands in place and skips the re-layout
struct alignas(64) FastMaps { CUtensorMap m[kMaxLevels]; CUtensorMap m48[kMaxLevels]; };
template <int WWIN_>
struct alignas(128) WarpSmem {
    static constexpr int WWIN = WWIN_;                    // max window side
    // 40-px windows are re-laid after the TMA load: 64-byte box rows put rows r and r+2 in the same banks (ncu: 2-way
    // conflicts in stage 1a, 3.8-way in stage 1b, 2.5-way on stage 2's byte loads); a 12-word pitch spreads eight
    // consecutive rows over distinct bank offsets.  The landing buffer is dead afterwards and becomes the candidate
    // list, so the block shrinks from 6400 to 6016 bytes and a ninth CTA fits an SM.
    static constexpr bool RELAY = WWIN_ <= 40;
    static constexpr int WS = RELAY ? 12 : 16;            // window row stride in words
    static constexpr int WZ = WWIN_ - 6;                  // max zone side
    static constexpr int SCS = WWIN_;                     // score-map row stride in bytes
    uint32_t land[RELAY ? WWIN_ : 1][16];                 // RELAY: TMA destination (64-byte box rows), later the list
    alignas(128) uint32_t win[WWIN][WS];                  // the window the stages read (the TMA destination otherwise)
    uint8_t sc[WZ + 2][SCS];
    unsigned short list_[RELAY ? 1 : WZ * WZ];
    unsigned long long bar;                               // mbarrier of the window load
    __device__ __forceinline__ unsigned short* list() { return RELAY ? reinterpret_cast<unsigned short*>(&land[0][0]) : list_; }
};
static_assert(sizeof(uint32_t) * 40 * 16 >= sizeof(unsigned short) * 34 * 34, "the landing buffer must hold the candidate list");

// per byte d > t: bit 7 of (d & 0x7F) + (0x7F - t) is "low seven bits > t" for t < 128 (then OR the byte's own bit 7),
// and with c = 0x7F - (t - 128) "low seven bits > t - 128" for t >= 128 (then AND it)
template <bool SMALL_T>
__device__ __forceinline__ uint32_t swar_gt_t(uint32_t d, uint32_t c) {
    const uint32_t m = (d & 0x7F7F7F7Fu) + c;       // per byte, no carry across bytes
    return (SMALL_T ? (m | d) : (m & d)) & 0x80808080u;
}

template <int WWIN_> struct RELAY_OF { static constexpr bool value = WarpSmem<WWIN_>::RELAY; };

template <int WWIN, bool SMALL_T>
__global__ void __launch_bounds__(kWarpsPerCta * 32, WWIN <= 40 ? 9 : 1)
fast_cells_warp_kernel(const __grid_constant__ FrameSetGeom g, const __grid_constant__ FastMaps maps,
                       const int4* __restrict__ celltab, int* __restrict__ cellc, uint32_t* __restrict__ cellk, int cellLo, int cellHi) {
    using Smem = WarpSmem<WWIN>;
    constexpr int WS = Smem::WS, SCS = Smem::SCS;
    static_assert(sizeof(S_sc_t<WWIN>) >= 4 * (WWIN - 6) * ((WWIN - 6 + 3) / 4 + 1), "task list must fit the score map");
    __shared__ Smem smem[kWarpsPerCta];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    Smem& S = smem[wid];
    const int frame = blockIdx.y + g.frame0, cell = cellLo + blockIdx.x * kWarpsPerCta + wid;
    if (cell >= cellHi) return;
    // host-built cell record: window origin / size, zone offset, level (api.cu: build_geometry)
    const int4 cd = __ldg(&celltab[cell]);
    const int l = cd.w & 15;
    const bool direct = RELAY_OF<WWIN>::value && (cd.w & 16);        // host: the window fits a 48-byte TMA box
    const uint32_t mdiv = (uint32_t)cd.w >> 8;                       // host: 65536 / nzc + 1
    const LevelGeom& L = g.lv[l];
    const int c = cell - L.cellBase;
    int* out_cnt = cellc + (long long)frame * g.ncells + cell;
    uint32_t* out = cellk + (long long)frame * g.cellSlots + L.cellSlotOff + (long long)c * L.cellCap;
    const int iniX = cd.x & 0xFFFF, iniY = cd.x >> 16;                          // :846-869
    const int ww = cd.y & 0xFFFF, wh = cd.y >> 16;
    const int offX = cd.z & 0xFFFF, offY = cd.z >> 16;                          // cell origin inside the border box
    const int maxX = iniX + ww;
    if (ww == 0) {                                                              // degenerate cell (:861-869)
        if (lane == 0) *out_cnt = 0;
        return;
    }
    const int zw = ww - 6, zh = wh - 6;
    // The window arrives as one TMA box: 64 bytes x WWIN rows starting at the 16-byte boundary left of the
    // cell (TMA wants 16-byte aligned row starts; rows past the cell are fetched and ignored).  Window byte
    // (r, x) below is relative to ax0: that boundary, or after the re-layout the 4-byte boundary left of the cell.
    constexpr bool RELAY = Smem::RELAY;
    const int tx0 = iniX & ~15;
    const int ax0 = RELAY && !direct ? iniX & ~3 : tx0;
    const uint32_t barA = (uint32_t)__cvta_generic_to_shared(&S.bar);
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barA) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barA), "r"(WWIN * (direct ? 48 : 64)) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"((uint32_t)__cvta_generic_to_shared(RELAY && !direct ? &S.land[0][0] : &S.win[0][0])),
                       "l"(reinterpret_cast<uint64_t>(direct ? &maps.m48[l] : &maps.m[l])), "r"(tx0), "r"(frame * L.h + iniY), "r"(barA) : "memory");
    }
    const int zx0 = iniX + 3 - ax0, zx1 = maxX - 3 - ax0;   // zone columns in window bytes
    const int zc0 = zx0 >> 2, nzc = ((zx1 - 1) >> 2) - zc0 + 1;
    // floor(a / nzc) for 0 <= a <= 32 and 1 <= nzc <= 16 is (a * mdiv) >> 16 with mdiv = 65536 / nzc + 1 from the host's
    // cell record (checked exhaustively there): no float reciprocal in the per-cell prologue
    const uint8_t* wb = reinterpret_cast<const uint8_t*>(&S.win[0][0]);       // byte (r, x) = wb[r*WS*4 + x]

    // stage 1a lane mapping: a lane keeps one zone word column and walks the rows rpi at a time
    // (3 rows per step whenever they fit: the ring offset is 3 too, so a lane's centre row becomes
    // its upper row one step later and one new word per step feeds both comparisons)
    const int rpi = nzc <= 10 ? 3 : (int)((32u * mdiv) >> 16);   // zone rows per warp step (nzc <= 12)
    const int rsub = (int)(((uint32_t)lane * mdiv) >> 16), cc1 = lane - rsub * nzc;
    const bool act = rsub < rpi;
    const uint32_t* wcol = &S.win[act ? rsub : 0][zc0 + cc1];    // idle lanes stay inside the block (their loads are unconditional)
    uint32_t zm = 0x80808080u;                                   // bytes of the lane's word inside the zone
    {
        const int xb = (zc0 + cc1) * 4;
        if (xb < zx0) zm &= 0xFFFFFFFFu << (8 * (zx0 - xb));
        if (xb + 4 > zx1) zm &= 0xFFFFFFFFu >> (8 * (xb + 4 - zx1));
    }

    __syncwarp();
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(barA), "r"(0) : "memory");
    }

    if (RELAY && !direct) {
        // two window rows per step: lanes 0-11 copy the 12 words of row r, lanes 12-23 those of row r+1 (conflict free
        // on both sides); the source starts at the word that holds column ax0
        const int half = lane >= 12 ? 1 : 0, cw = lane - 12 * half;
        const uint32_t* srcw = &S.land[half][((iniX & 15) >> 2) + cw];
        uint32_t* dstw = &S.win[half][cw];
        // all WWIN rows (the box always brings them), fully unrolled: immediate offsets, no loop bookkeeping
        if (lane < 24) {
#pragma unroll
            for (int k = 0; k < WWIN / 2; k++) dstw[k * 2 * WS] = srcw[k * 32];
        }
        __syncwarp();
    }
    int total = 0;
    for (int pass = 0; pass < 2; pass++) {
        const int t = pass == 0 ? g.iniTh : g.minTh;
        // SMALL_T: both thresholds are below 128 (every configuration the reference ships); otherwise decided per pass
        const bool small_t = SMALL_T || t < 128;
        const uint32_t cmpc = (uint32_t)(small_t ? 0x7F - t : 0x7F - (t - 128)) * 0x01010101u;
        auto swar_gt = [&](uint32_t dd, uint32_t cc, bool) { return SMALL_T ? swar_gt_t<true>(dd, cc) : (small_t ? swar_gt_t<true>(dd, cc) : swar_gt_t<false>(dd, cc)); };
        // ---- stage 1a: vertical pair (0,8) on every zone word; surviving words are compacted
        // (ballot order) into a task list that borrows the score map's storage
        uint32_t* tasks = reinterpret_cast<uint32_t*>(&S.sc[0][0]);
        int ntask = 0;
        if (rpi == 3) {
            // loads are unconditional (rows past the zone and the idle lanes read other parts of this warp's shared
            // block and are masked out), so the loop body has no divergent branch.  Steps whose three rows all lie in
            // the zone need no row test; at most one tail step does.
            uint32_t ce = wcol[3 * WS];
            uint32_t g_up = swar_gt(__vabsdiffu4(wcol[0], ce), cmpc, small_t);
            const uint32_t zma = act ? zm : 0u;
            uint32_t tagr = (uint32_t)rsub | ((uint32_t)cc1 << 8);          // tag + r0 of the current step
            const uint32_t lt = (1u << lane) - 1u;
            uint32_t* tp = tasks;                                            // tasks + ntask
            const uint32_t* wp = wcol + 6 * WS;                              // (r0 + 6) * WS
            auto step1a = [&](uint32_t rowmask) {
                const uint32_t dn = *wp;
                const uint32_t g_dn = swar_gt(__vabsdiffu4(dn, ce), cmpc, small_t);
                const uint32_t gmask = (g_up | g_dn) & rowmask;
                ce = dn; g_up = g_dn;                            // |centre - below| is |above - centre| of the next step
                const uint32_t bal = __ballot_sync(0xFFFFFFFFu, gmask != 0);
                if (gmask) tp[__popc(bal & lt)] = gmask | tagr;
                tp += __popc(bal);
                tagr += 3; wp += 3 * WS;
            };
            int r0 = 0;
            for (; r0 + 3 <= zh; r0 += 3) step1a(zma);
            if (r0 < zh) step1a(r0 + rsub < zh ? zma : 0u);
            ntask = (int)(tp - tasks);
        } else {
            for (int r0 = 0; r0 < zh; r0 += rpi) {
                const int r = r0 + rsub;
                uint32_t gmask = 0;
                if (act && r < zh) {
                    const uint32_t* pw = wcol + r0 * WS;             // (window row r, word column wc)
                    const uint32_t v = pw[3 * WS];
                    gmask = (swar_gt(__vabsdiffu4(pw[6 * WS], v), cmpc, small_t) |
                             swar_gt(__vabsdiffu4(pw[0], v), cmpc, small_t)) & zm;
                }
                const uint32_t bal = __ballot_sync(0xFFFFFFFFu, gmask != 0);
                if (gmask) tasks[ntask + __popc(bal & ((1u << lane) - 1u))] = gmask | (uint32_t)r | ((uint32_t)cc1 << 8);
                ntask += __popc(bal);
            }
        }
        __syncwarp();
        // ---- stage 1b: pairs (4,12), (2,10), (6,14) on the surviving words, all lanes busy.
        // Every list from here on is built with warp scans in task order, which is row-major:
        // the emitted corners come out in cv::FAST's order without sorting or atomics.
        int ncand = 0;
        for (int base = 0; base < ntask; base += 32) {
            const int i = base + lane;
            uint32_t gmask = 0;
            int r = 0, xz = 0;
            if (i < ntask) {
                const uint32_t e = tasks[i];
                gmask = e & 0x80808080u;
                r = e & 0x7F;
                const int cc = (e >> 8) & 0x7F;
                const int y = r + 3, wc = zc0 + cc;
                xz = (zc0 + cc) * 4 - zx0;                        // zone column of the word's byte 0
                const uint32_t v = S.win[y][wc];
                const uint32_t wl = S.win[y][wc - 1], wr = S.win[y][wc + 1];
                gmask &= swar_gt(__vabsdiffu4(__funnelshift_r(v, wr, 24), v), cmpc, small_t) |
                         swar_gt(__vabsdiffu4(__funnelshift_r(wl, v, 8), v), cmpc, small_t);
                const uint32_t r2 = __funnelshift_r(S.win[y + 2][wc], S.win[y + 2][wc + 1], 16);
                const uint32_t r10 = __funnelshift_r(S.win[y - 2][wc - 1], S.win[y - 2][wc], 16);
                gmask &= swar_gt(__vabsdiffu4(r2, v), cmpc, small_t) | swar_gt(__vabsdiffu4(r10, v), cmpc, small_t);
                const uint32_t r6 = __funnelshift_r(S.win[y - 2][wc], S.win[y - 2][wc + 1], 16);
                const uint32_t r14 = __funnelshift_r(S.win[y + 2][wc - 1], S.win[y + 2][wc], 16);
                gmask &= swar_gt(__vabsdiffu4(r6, v), cmpc, small_t) | swar_gt(__vabsdiffu4(r14, v), cmpc, small_t);
            }
            const int k = __popc(gmask);
            int inc = k;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xFFFFFFFFu, inc, o);
                if (lane >= o) inc += u;
            }
            int pos = ncand + inc - k;
#pragma unroll
            for (int b = 0; b < 4; b++)
                if (gmask & (0x80u << (8 * b))) S.list()[pos++] = (unsigned short)((r << 8) | (xz + b));
            ncand += __shfl_sync(0xFFFFFFFFu, inc, 31);
        }
        __syncwarp();
        // the task list is consumed: zero the score map
        // (the whole map, 16 bytes per lane and step: three or five steps)
        static_assert(sizeof(S.sc) % 16 == 0 && offsetof(Smem, sc) % 16 == 0, "score map is cleared with 16-byte stores");
#pragma unroll
        for (int i = 0; i < (int)(sizeof(S.sc) / 16 + 31) / 32; i++)
            if (i * 32 + lane < (int)(sizeof(S.sc) / 16)) reinterpret_cast<uint4*>(&S.sc[0][0])[i * 32 + lane] = make_uint4(0u, 0u, 0u, 0u);
        __syncwarp();
        // ---- stage 2: exact test + strength on the survivors, compacted in place
        int ncorn = 0;
        for (int base = 0; base < ncand; base += 32) {
            const int i = base + lane;
            int e = 0, s = 0;
            if (i < ncand) {
                e = S.list()[i];
                const int r = e >> 8, cc = e & 0xFF;
                s = corner_strength<false>(wb + (r + 3) * (WS * 4) + zx0 + cc, WS * 4, t);
                if (s) S.sc[r + 1][cc + 1] = (uint8_t)s;
            }
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, s != 0);
            __syncwarp();                                 // list reads above before the in-place writes
            if (s) S.list()[ncorn + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)e;
            ncorn += __popc(bal);
        }
        __syncwarp();
        // ---- stage 3: 3x3 strict maximum inside the zone; survivors leave in list order
        total = 0;
        for (int base = 0; base < ncorn; base += 32) {
            const int i = base + lane;
            bool keep = false;
            int r = 0, cc = 0, s = 0;
            if (i < ncorn) {
                const int e = S.list()[i];
                r = e >> 8, cc = e & 0xFF;
                s = S.sc[r + 1][cc + 1];
                keep = s > S.sc[r][cc] && s > S.sc[r][cc + 1] && s > S.sc[r][cc + 2] &&
                       s > S.sc[r + 1][cc] && s > S.sc[r + 1][cc + 2] && s > S.sc[r + 2][cc] &&
                       s > S.sc[r + 2][cc + 1] && s > S.sc[r + 2][cc + 2];
            }
            const uint32_t bal = __ballot_sync(0xFFFFFFFFu, keep);
            const int pos = total + __popc(bal & ((1u << lane) - 1u));
            if (keep && pos < L.cellCap) out[pos] = pack_kp(cc + 3 + offX, r + 3 + offY, s - 1);
            total += __popc(bal);
        }
        if (total > 0) break;
        __syncwarp();
    }
    if (lane == 0) *out_cnt = min(total, L.cellCap);
}

}  // namespace

// window rows of the warp kernel that takes level l (0: generic kernel) — also the TMA box height
int fast_window_rows(const LevelGeom& L) {
    if (!L.nCols) return 40;
    const int side = (L.wCell > L.hCell ? L.wCell : L.hCell) + 6;
    return side > kWarpWinMax ? 0 : side > 40 ? 48 : 40;
}

int launch_fast(const FrameSetGeom& g, const DevBufs& d, cudaStream_t s, int l0, int l1) {
    int launches = 0;
    if (l1 < 0 || l1 > g.nlevels) l1 = g.nlevels;
    if (l0 < 0) l0 = 0;
    FastMaps maps;
    std::memcpy(maps.m, d.fast_tmaps, sizeof(CUtensorMap) * g.nlevels);
    std::memcpy(maps.m48, d.fast_tmaps48, sizeof(CUtensorMap) * g.nlevels);
    // runs of consecutive levels that share a kernel variant: 0 = generic, 40 / 48 = warp layouts
    bool generic = false;
    auto variant = [&](int l) {
        if (!g.lv[l].nCols) return -1;
        const int side = (g.lv[l].wCell > g.lv[l].hCell ? g.lv[l].wCell : g.lv[l].hCell) + 6;
        return side > kWarpWinMax ? 0 : side > 40 ? 48 : 40;
    };
    auto cell_end = [&](int l) { return l < g.nlevels ? g.lv[l].cellBase : g.ncells; };
    for (int l = l0; l < l1;) {
        const int v = variant(l);
        int e = l + 1;
        while (e < l1 && (variant(e) == v || variant(e) < 0)) e++;
        const int lo = g.lv[l].cellBase, hi = cell_end(e);
        if (v == 0) generic = true;
        if (v > 0 && hi > lo) {
            const dim3 grid((hi - lo + kWarpsPerCta - 1) / kWarpsPerCta, g.batch);
            const bool smallT = g.iniTh < 128 && g.minTh < 128;
            auto kern = v == 40 ? (smallT ? fast_cells_warp_kernel<40, true> : fast_cells_warp_kernel<40, false>)
                                : (smallT ? fast_cells_warp_kernel<48, true> : fast_cells_warp_kernel<48, false>);
            kern<<<grid, kWarpsPerCta * 32, 0, s>>>(g, maps, d.celltab, d.cellc, d.cellk, lo, hi);
            launches++;
        }
        l = e;
    }
    if (generic) {
        // the generic kernel skips the cells the warp kernels took
        const int lo = g.lv[l0].cellBase, hi = cell_end(l1);
        if (hi > lo) {
            fast_cells_kernel<<<dim3(hi - lo, g.batch), kThreads, 0, s>>>(g, d.pyr, d.cellc, d.cellk, lo);
            launches++;
        }
    }
    return launches;
}

}  // namespace orbfe
