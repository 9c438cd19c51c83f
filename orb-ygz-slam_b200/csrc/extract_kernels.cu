// extract_kernels.cu — sm_100a kernels of the ORB extractor hot path.
// Every kernel is batched over frames (blockIdx.y = frame) so that one launch covers a whole
// batch of same-sized frames.  Integer arithmetic reproduces OpenCV 4.13's fixed-point paths bit
// for bit; float arithmetic is written with explicit round-to-nearest intrinsics so nvcc cannot
// contract it into FMAs the reference's build (-ffp-contract=off) does not execute.
#include "ygz_common.cuh"
#include "ygz_dso.cuh"
#include "orb_pattern.inc"

namespace ygz {

// <<<>>> or, with B.pdl, cudaLaunchKernelEx with programmatic stream serialization (the kernel may start while its predecessor
// in the stream still runs; pdlEnter() in the kernel orders the data)
template <class... KArgs, class... Args>
static void launchChain(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, args...);
}

// =================================================================================================
// K1 — pyramid level by cv::resize(INTER_LINEAR), 8UC1 (src/ORBextractor.cc:1139).
// OpenCV's HResizeLinear/VResizeLinear fixed point: 11-bit coefficients, horizontal pass in int32,
// vertical pass ((b0*(r0>>4))>>16 + (b1*(r1>>4))>>16 + 2) >> 2.  The coefficient tables are built on
// the host (extractor.cu, makeResizeTables) with the double/float recipe of resize.cpp, so the
// kernel is integer-only.  One thread produces 4 horizontally adjacent pixels (one 32-bit store).
// Table entry X[dx] = {x0 | x1<<16, a0 | a1<<16}; Y[dy] = {y0 | y1<<16, b0 | b1<<16}.
constexpr int kResizeBlockY = 4;   // block = 32 x kResizeBlockY threads
constexpr int kResizeRows = 8;   // destination rows per thread

template <bool PDL> __global__ void __launch_bounds__(256) k_resize_generic(const __grid_constant__ FrameGeom G, const DevBuffers B, int level) {
    if (PDL) pdlEnter(B);
    const LevelGeom &D = G.lv[level];
    const LevelGeom &S = G.lv[level - 1];
    const int f = blockIdx.z;
    const int dx0 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int dyBase = (blockIdx.y * kResizeBlockY + threadIdx.y) * kResizeRows;
    if (dx0 >= D.pitch || dyBase >= D.h) return;
    const uint8_t *src = levelPtr(G, B, f, level - 1);
    uint8_t *dst = B.pyr + (long long)f * G.slabBytes + D.off + dx0;
    const int sp = S.pitch;
    if (D.area2x) {  // ResizeAreaFast 2x2 mean
        for (int k = 0; k < kResizeRows; k++) {
            const int dy = dyBase + k;
            if (dy >= D.h) break;
            const uint8_t *s0 = src + (long long)(2 * dy) * sp, *s1 = s0 + sp;
            uint32_t out = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int dx = min(dx0 + i, D.w - 1);
                const int v = (s0[2 * dx] + s0[2 * dx + 1] + s1[2 * dx] + s1[2 * dx + 1] + 2) >> 2;
                out |= (uint32_t)v << (8 * i);
            }
            *reinterpret_cast<uint32_t *>(dst + (long long)dy * D.pitch) = out;
        }
        return;
    }
    // column taps of the 4 pixels, once per thread (the table is padded to the pitch)
    int x0[4], x1[4], a0[4], a1[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int2 tx = __ldg(B.tabs + D.tabX + dx0 + i);
        x0[i] = tx.x & 0xffff; x1[i] = (uint32_t)tx.x >> 16;
        a0[i] = (short)(tx.y & 0xffff); a1[i] = (short)((uint32_t)tx.y >> 16);
    }
    // rows: the horizontal pass of source row y1(dy) is reused when it is y0(dy + 1) (always, away from the edges)
    int prevY = -1, hPrev[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < kResizeRows; k++) {
        const int dy = dyBase + k;
        if (dy < D.h) {
            const int2 ty = __ldg(B.tabs + D.tabY + dy);
            const int y0 = ty.x & 0xffff, y1 = (uint32_t)ty.x >> 16;
            const int b0 = (short)(ty.y & 0xffff), b1 = (short)((uint32_t)ty.y >> 16);
            int h0[4], h1[4];
            if (y0 == prevY) {
#pragma unroll
                for (int i = 0; i < 4; i++) h0[i] = hPrev[i];
            } else {
                const uint8_t *s0 = src + (long long)y0 * sp;
#pragma unroll
                for (int i = 0; i < 4; i++) h0[i] = s0[x0[i]] * a0[i] + s0[x1[i]] * a1[i];
            }
            if (y1 == y0) {
#pragma unroll
                for (int i = 0; i < 4; i++) h1[i] = h0[i];
            } else {
                const uint8_t *s1 = src + (long long)y1 * sp;
#pragma unroll
                for (int i = 0; i < 4; i++) h1[i] = s1[x0[i]] * a0[i] + s1[x1[i]] * a1[i];
            }
            uint32_t out = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                int v = (((b0 * (h0[i] >> 4)) >> 16) + ((b1 * (h1[i] >> 4)) >> 16) + 2) >> 2;
                v = min(max(v, 0), 255);
                out |= (uint32_t)v << (8 * i);
                hPrev[i] = h1[i];
            }
            prevY = y1;
            *reinterpret_cast<uint32_t *>(dst + (long long)dy * D.pitch) = out;  // pitch is a multiple of 64: padding is scratch
        }
    }
}

// Word-based variant used whenever the level qualifies (LevelGeom::fastResize, built on the host: all weights
// >= 0, the second tap is the next column or has weight 0, the 4 pixels of a thread span <= 7 source columns;
// plus 4-byte aligned source rows).  Per source row a thread loads the three aligned words that hold its
// columns, funnel-shifts them to an 8-byte window that starts at its first column, and gets each pixel's
// horizontal sum from ONE byte permutation (the two neighbouring source bytes) and ONE DP2A against the packed
// 16-bit weights (a0 | a1 << 16) exactly as the table stores them.  The vertical pass uses the high half of
// the 32-bit products ((b * (h >> 4)) >> 16); the saturate_cast of the reference is the identity here
// (weights sum to 2048, so the result is <= 255).
template <bool PDL> __global__ void __launch_bounds__(256) k_resize(const __grid_constant__ FrameGeom G, const DevBuffers B, int level) {
    if (PDL) pdlEnter(B);
    const LevelGeom &D = G.lv[level];
    const LevelGeom &S = G.lv[level - 1];
    const int f = blockIdx.z;
    const int dx0 = (blockIdx.x * 32 + threadIdx.x) * 4;
    const int dyBase = (blockIdx.y * kResizeBlockY + threadIdx.y) * kResizeRows;
    if (dx0 >= D.pitch || dyBase >= D.h) return;
    const uint8_t *src = levelPtr(G, B, f, level - 1);
    uint8_t *dst = B.pyr + (long long)f * G.slabBytes + D.off + dx0;
    const int sp = S.pitch;
    // column taps, once per thread (the table is padded to the pitch)
    uint32_t coef[4], sel[4];
    int base;
    {
        const int4 *tp = reinterpret_cast<const int4 *>(B.tabs + D.tabX + dx0);   // 2 entries per 16-byte load
        const int4 t01 = __ldg(tp), t23 = __ldg(tp + 1);
        base = t01.x & 0xffff;
        const int d1 = (t01.z & 0xffff) - base, d2 = (t23.x & 0xffff) - base, d3 = (t23.z & 0xffff) - base;
        coef[0] = t01.y; coef[1] = t01.w; coef[2] = t23.y; coef[3] = t23.w;
        sel[0] = 0x10u; sel[1] = d1 | ((d1 + 1) << 4); sel[2] = d2 | ((d2 + 1) << 4); sel[3] = d3 | ((d3 + 1) << 4);
    }
    const int maxWord = (sp >> 2) - 1, wb = base >> 2;
    const int i0 = min(wb, maxWord), i1 = min(wb + 1, maxWord), i2 = min(wb + 2, maxWord);
    const int sh0 = 8 * (base & 3);
    auto hpass = [&](int y, uint32_t h[4]) {
        const uint32_t *row = reinterpret_cast<const uint32_t *>(src + (long long)y * sp);
        const uint32_t w0 = __ldg(row + i0), w1 = __ldg(row + i1), w2 = __ldg(row + i2);
        const uint32_t v0 = __funnelshift_r(w0, w1, sh0), v1 = __funnelshift_r(w1, w2, sh0);
#pragma unroll
        for (int i = 0; i < 4; i++) h[i] = __dp2a_lo(coef[i], __byte_perm(v0, v1, sel[i]), 0u) >> 4;
    };
    // rows: the horizontal pass of source row y1(dy) is reused when it is y0(dy + 1) (warp-uniform branches)
    int prevY = -1;
    uint32_t hPrev[4] = {0, 0, 0, 0};
#pragma unroll
    for (int k = 0; k < kResizeRows; k++) {
        const int dy = dyBase + k;
        if (dy < D.h) {
            const int2 ty = __ldg(B.tabs + D.tabY + dy);
            const int y0 = ty.x & 0xffff, y1 = (uint32_t)ty.x >> 16;
            const uint32_t b0 = (uint32_t)(ty.y & 0xffff) << 16, b1 = (uint32_t)ty.y & 0xffff0000u;
            uint32_t h0[4], h1[4];
            if (y0 == prevY) {
#pragma unroll
                for (int i = 0; i < 4; i++) h0[i] = hPrev[i];
            } else {
                hpass(y0, h0);
            }
            if (y1 == y0) {
#pragma unroll
                for (int i = 0; i < 4; i++) h1[i] = h0[i];
            } else {
                hpass(y1, h1);
            }
            uint32_t out = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const uint32_t v = (__umulhi(b0, h0[i]) + __umulhi(b1, h1[i]) + 2u) >> 2;
                out |= v << (8 * i);
                hPrev[i] = h1[i];
            }
            prevY = y1;
            *reinterpret_cast<uint32_t *>(dst + (long long)dy * D.pitch) = out;  // pitch is a multiple of 64: padding is scratch
        }
    }
}

// Band variant (the one batches run): CTA = 64 destination columns x nSub sub-tiles of kRzTH destination rows; 256 threads.
// Per sub-tile the horizontal pass of EVERY source row it touches (rows y0(first dy) .. y1(last dy), about 1.2 x kRzTH + 2 of
// them) goes to shared memory once, as four 32-bit sums per column group (thread = column group x row slot); the vertical pass
// then picks its two rows by index (thread = two column groups x row slot).  Against k_resize (register-resident row reuse)
// this has no data-dependent branches, no table decode in the horizontal pass (the host pre-digests the X table per column
// group: weights, byte selectors, window offset, shift) and one address update per source row.
// A column group whose 8-byte window would cross the end of the source row has it slid back on the host (offset - delta,
// selectors + delta), so the three word loads stay inside the row (the third one is folded onto the second when the slid window
// ends exactly at the row's last word).
constexpr int kRzTH = 64, kRzThreads = 256, kRzSlots = 16, kRzMaxBand = 160;

// STAGED = the source band goes global -> shared as 16-byte cp.async chunks (7 per row: the strip's source columns from
// xs = first window & ~15, kRzRawPitch = 112 bytes; levels that need more — scale factors above ~1.3 — take the LDG form) — no registers, the whole band in flight at once — and the horizontal pass reads it with
// LDS; the band of sub-tile s + 1 is requested right after the horizontal pass of sub-tile s, so it travels while the vertical
// pass of s runs.  Needs 16-byte aligned source rows (base and pitch); otherwise the horizontal pass loads its words with LDG.
constexpr int kRzRawPitch = 112;   // staged bytes per source row of a strip: enough for the 1.2x pyramid (LevelGeom::rawPitch says what a level needs)
__device__ __forceinline__ void cpAsync16(void *smemDst, const void *gmemSrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smemDst)), "l"(gmemSrc));
}
__device__ __forceinline__ void rzStageBand(uint8_t *raw, int rawPitch, const uint8_t *srcStrip, int sp, int xs, int yFirst, int yLast, int tid) {
    // thread = (chunk column tid & 15, row tid >> 4 + 16 i)
    const int c = tid & 15;
    if (16 * c < rawPitch && xs + 16 * c + 16 <= sp) {
        const uint8_t *g = srcStrip + (long long)(yFirst + (tid >> 4)) * sp + 16 * c;
        uint8_t *d = raw + (tid >> 4) * rawPitch + 16 * c;
        for (int y = yFirst + (tid >> 4); y <= yLast; y += kRzThreads / 16, g += (long long)(kRzThreads / 16) * sp, d += (kRzThreads / 16) * rawPitch)
            cpAsync16(d, g);
    }
}

template <bool PDL, bool STAGED> __global__ void __launch_bounds__(kRzThreads, 6) k_resize_band(const __grid_constant__ FrameGeom G, const DevBuffers B, int level, int nSub) {
    if (PDL) pdlEnter(B);
    extern __shared__ __align__(16) uint2 rzH[];   // [band row][16 column groups]: four 16-bit sums (h >> 4 <= 32640) per group; STAGED: the raw band behind it
    const LevelGeom &D = G.lv[level];
    const LevelGeom &S = G.lv[level - 1];
    const int f = blockIdx.z, tid = threadIdx.x;
    const int cg = tid & 15, slot = tid >> 4;              // horizontal pass mapping
    const int cp = tid & 7, vslot = tid >> 3;              // vertical pass mapping: column groups cp and cp + 8
    const int sp = S.pitch;
    const int2 *tY = B.tabs + D.tabY;
    // pre-digested X table: per column group {coef[4]}, {sel[4]}, {window word offset, shift, third-word offset, strip start xs}
    const int4 *tg = reinterpret_cast<const int4 *>(B.tabs + D.tabXG) + 3 * (blockIdx.x * 16 + cg);
    const int4 coef = __ldg(tg), sel = __ldg(tg + 1), misc = __ldg(tg + 2);
    const uint8_t *srcLevel = levelPtr(G, B, f, level - 1);
    const uint8_t *srcCol = srcLevel + misc.x;
    const long long step = (long long)kRzSlots * sp;
    uint8_t *dst = B.pyr + (long long)f * G.slabBytes + D.off + (blockIdx.x * 16 + cp) * 4;
    const uint2 *hc = rzH + cp;
    uint8_t *raw = reinterpret_cast<uint8_t *>(rzH + D.bandRows * 16);
    const int dy0 = blockIdx.y * nSub * kRzTH;
    if (dy0 >= D.h) return;
    int yFirst = __ldg(tY + dy0).x & 0xffff, yLast = (int)((uint32_t)__ldg(tY + min(dy0 + kRzTH - 1, D.h - 1)).x >> 16);
    constexpr int rawPitch = kRzRawPitch;
    if (STAGED) rzStageBand(raw, rawPitch, srcLevel + misc.w, sp, misc.w, yFirst, yLast, tid);
    for (int sub = 0; sub < nSub; sub++) {
        const int dyBase = dy0 + sub * kRzTH;
        if (dyBase >= D.h) break;
        const int dyLast = min(dyBase + kRzTH - 1, D.h - 1);
        if (STAGED) asm volatile("cp.async.wait_all;" ::: "memory");
        if (STAGED || sub > 0) __syncthreads();            // the band has landed; the previous sub-tile's vertical pass is done with rzH
        if (STAGED) {
            const uint8_t *p = raw + slot * rawPitch + (misc.x - misc.w);
            uint2 *hd = rzH + slot * 16 + cg;
#pragma unroll 2
            for (int y = yFirst + slot; y <= yLast; y += kRzSlots, p += kRzSlots * rawPitch, hd += kRzSlots * 16) {
                const uint32_t w0 = *reinterpret_cast<const uint32_t *>(p), w1 = *reinterpret_cast<const uint32_t *>(p + 4);
                const uint32_t w2 = *reinterpret_cast<const uint32_t *>(p + misc.z);
                const uint32_t v0 = __funnelshift_r(w0, w1, misc.y), v1 = __funnelshift_r(w1, w2, misc.y);
                const uint32_t h0 = __dp2a_lo((uint32_t)coef.x, __byte_perm(v0, v1, sel.x), 0u) >> 4;
                const uint32_t h1 = __dp2a_lo((uint32_t)coef.y, __byte_perm(v0, v1, sel.y), 0u) >> 4;
                const uint32_t h2 = __dp2a_lo((uint32_t)coef.z, __byte_perm(v0, v1, sel.z), 0u) >> 4;
                const uint32_t h3 = __dp2a_lo((uint32_t)coef.w, __byte_perm(v0, v1, sel.w), 0u) >> 4;
                *hd = make_uint2(h1 * 65536u + h0, h3 * 65536u + h2);
            }
        } else {
            const uint8_t *p = srcCol + (long long)(yFirst + slot) * sp;
            uint2 *hd = rzH + slot * 16 + cg;
#pragma unroll 2
            for (int y = yFirst + slot; y <= yLast; y += kRzSlots, p += step, hd += kRzSlots * 16) {
                const uint32_t w0 = __ldg(reinterpret_cast<const uint32_t *>(p)), w1 = __ldg(reinterpret_cast<const uint32_t *>(p + 4));
                const uint32_t w2 = __ldg(reinterpret_cast<const uint32_t *>(p + misc.z));
                const uint32_t v0 = __funnelshift_r(w0, w1, misc.y), v1 = __funnelshift_r(w1, w2, misc.y);
                const uint32_t h0 = __dp2a_lo((uint32_t)coef.x, __byte_perm(v0, v1, sel.x), 0u) >> 4;
                const uint32_t h1 = __dp2a_lo((uint32_t)coef.y, __byte_perm(v0, v1, sel.y), 0u) >> 4;
                const uint32_t h2 = __dp2a_lo((uint32_t)coef.z, __byte_perm(v0, v1, sel.z), 0u) >> 4;
                const uint32_t h3 = __dp2a_lo((uint32_t)coef.w, __byte_perm(v0, v1, sel.w), 0u) >> 4;
                *hd = make_uint2(h1 * 65536u + h0, h3 * 65536u + h2);
            }
        }
        __syncthreads();
        const int yF = yFirst;
        if (sub + 1 < nSub && dyBase + kRzTH < D.h) {       // next sub-tile's band: known now, requested now (STAGED), used after the vertical pass
            yFirst = __ldg(tY + dyBase + kRzTH).x & 0xffff;
            yLast = (int)((uint32_t)__ldg(tY + min(dyBase + 2 * kRzTH - 1, D.h - 1)).x >> 16);
            if (STAGED) rzStageBand(raw, rawPitch, srcLevel + misc.w, sp, misc.w, yFirst, yLast, tid);
        }
#pragma unroll 2
        for (int dy = dyBase + vslot; dy <= dyLast; dy += kRzThreads / 8) {
            const int2 ty = __ldg(tY + dy);
            const int y0 = (ty.x & 0xffff) - yF, y1 = (int)((uint32_t)ty.x >> 16) - yF;
            const uint32_t b0 = (uint32_t)ty.y << 16, b1 = (uint32_t)ty.y & 0xffff0000u;
            uint8_t *d = dst + (long long)dy * D.pitch;
#pragma unroll
            for (int half = 0; half < 2; half++) {
                const uint2 a = hc[y0 * 16 + 8 * half], b = hc[y1 * 16 + 8 * half];
                const uint32_t r0 = (__umulhi(b0, a.x & 0xffffu) + __umulhi(b1, b.x & 0xffffu) + 2u) >> 2, r1 = (__umulhi(b0, a.x >> 16) + __umulhi(b1, b.x >> 16) + 2u) >> 2;
                const uint32_t r2 = (__umulhi(b0, a.y & 0xffffu) + __umulhi(b1, b.y & 0xffffu) + 2u) >> 2, r3 = (__umulhi(b0, a.y >> 16) + __umulhi(b1, b.y >> 16) + 2u) >> 2;
                *reinterpret_cast<uint32_t *>(d + 32 * half) = (r3 * 256u + r2) * 65536u + (r1 * 256u + r0);  // padding columns are scratch
            }
        }
    }
}

int launchResizeLevel(const FrameGeom &G, const DevBuffers &B, int level, int nFrames, cudaStream_t s) {
    const LevelGeom &D = G.lv[level];
    const LevelGeom &S = G.lv[level - 1];
    dim3 block(32, kResizeBlockY), grid((D.pitch / 4 + 31) / 32, (D.h + kResizeBlockY * kResizeRows - 1) / (kResizeBlockY * kResizeRows), nFrames);
    // level 0 is caller memory: the word path needs 4-byte aligned rows there (levels >= 1 always are)
    const bool rowsAligned = (S.pitch & 3) == 0 &&
                             (level > 1 || ((reinterpret_cast<uintptr_t>(B.in) & 3) == 0 && (B.inFrameStride & 3) == 0));
    const bool word = D.fastResize && !D.area2x && rowsAligned;
    if (word && D.bandRows > 0 && D.bandRows <= kRzMaxBand) {
        // sub-tiles per CTA: a whole column strip for batches (the per-thread set-up is paid once per strip), one for small
        // batches (more CTAs)
        const int tilesY = (D.h + kRzTH - 1) / kRzTH;
        const int nSub = (long long)(D.pitch / 64) * nFrames >= 2048 ? tilesY : ((long long)(D.pitch / 64) * nFrames * ((tilesY + 1) / 2) >= 1024 ? 2 : 1);
        const dim3 bgrid(D.pitch / 64, (tilesY + nSub - 1) / nSub, nFrames);
        // cp.async staging of the source band needs 16-byte aligned source rows
        const uint8_t *srcBase = level == 1 ? B.in : B.pyr + S.off;
        const long long srcFrame = level == 1 ? B.inFrameStride : G.slabBytes;
        const bool staged = D.rawPitch > 0 && D.rawPitch <= kRzRawPitch && (S.pitch & 15) == 0 && (reinterpret_cast<uintptr_t>(srcBase) & 15) == 0 && (srcFrame & 15) == 0;
        const size_t smem = (size_t)D.bandRows * 16 * sizeof(uint2) + (staged ? (size_t)D.bandRows * kRzRawPitch : 0);
        if (staged) {
            if (B.pdl) launchChain(true, k_resize_band<true, true>, bgrid, dim3(kRzThreads), smem, s, G, B, level, nSub);
            else k_resize_band<false, true><<<bgrid, kRzThreads, smem, s>>>(G, B, level, nSub);
        } else {
            if (B.pdl) launchChain(true, k_resize_band<true, false>, bgrid, dim3(kRzThreads), smem, s, G, B, level, nSub);
            else k_resize_band<false, false><<<bgrid, kRzThreads, smem, s>>>(G, B, level, nSub);
        }
        return 1;
    }
    if (B.pdl) { if (word) launchChain(true, k_resize<true>, grid, block, 0, s, G, B, level); else launchChain(true, k_resize_generic<true>, grid, block, 0, s, G, B, level); }
    else if (word) k_resize<false><<<grid, block, 0, s>>>(G, B, level);
    else k_resize_generic<false><<<grid, block, 0, s>>>(G, B, level);
    return 1;
}

// =================================================================================================
// K2 — per-cell FAST-9/16 + 3x3 NMS + threshold fallback (src/ORBextractor.cc:747-781 calling
// cv::FAST twice per cell).  One CTA per cell.  The corner measure m (max over the 16 arcs of 9
// contiguous circle pixels of min(v-p) resp. min(p-v)) does not depend on the threshold: a pixel
// is a corner at threshold t iff m > t and its OpenCV score is m-1.  So the CTA computes m once for
// the cell's detection area (the cell minus a 3-px frame), runs the in-cell NMS at iniThFAST, and
// only if nothing survives runs it again at minThFAST — exactly the reference's "if empty, retry".
// NMS neighbours outside the detection area count as 0, as in cv::FAST on the cell sub-image.
// Survivors are appended (unordered) to the level's candidate array with an order key
// (cell, y-in-cell, x-in-cell) that encodes the reference's vToDistributeKeys order, which is all
// the quadtree needs from the order (its only order dependence is "first wins" on response ties).
// Tried in round 2 and dropped (profiles/r2_k_fast_experiments.md): a persistent variant fed by TMA (cp.async.bulk.tensor.3d, two tile
// buffers, mbarrier completion) — bit-exact, 2.27-2.50 ms against 1.91 ms: the 16-byte alignment TMA wants of the x coordinate costs
// a 64-byte tile pitch, i.e. shared memory for 16-21 CTAs per SM instead of 31, and a CTA then serialises its cells' barrier chains.
constexpr int kFastThreads = kFastCtaThreads;
constexpr int kFastWarps = kFastThreads / 32;
// image / score tile pitch = template parameter TP (kFastTilePitchNarrow or kFastTilePitch)

// The 16 ring differences are kept as packed 16-bit pairs E = (d + 256) | (256 - d) << 16 with
// d = v - p, both halves positive, so ONE VIMNMX3.U16x2 handles the "darker" arcs (low half) and
// the "brighter" arcs (high half) at once.  E = base + p * 0xFFFF with base = packed (v + 256, 256 - v):
// no borrow crosses the halves because v + 256 - p >= 1.
// Shared-memory accessors on 32-bit shared-window addresses.  With generic pointers the compiler
// re-derives the CTA's shared window base (S2UR CgaCtaId + ULEA) at every use inside the pixel loop;
// taking the address once with __cvta_generic_to_shared keeps the loop to the loads themselves.
__device__ __forceinline__ unsigned smemAddr(const void *p) {
    unsigned a = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("mov.u32 %0, %0;" : "+r"(a));   // opaque: stops the rematerialisation of the window base at every use
    return a;
}
template <int IMM> __device__ __forceinline__ unsigned ldsU8(unsigned a) {
    unsigned v;
    asm volatile("ld.shared.u8 %0, [%1+%2];" : "=r"(v) : "r"(a), "n"(IMM));
    return v;
}
__device__ __forceinline__ unsigned ldsU16(unsigned a) {
    unsigned v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void stsU8(unsigned a, unsigned v) { asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(v)); }
__device__ __forceinline__ void stsU16(unsigned a, unsigned v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "r"(v)); }

__device__ __forceinline__ unsigned fastBase(int v) { return (unsigned)(v + 256) * 0xFFFF0001u + (512u << 16); }
__device__ __forceinline__ unsigned fastE(int p, unsigned base) { return (unsigned)p * 0xFFFFu + base; }

// Compass pretest: an arc of 9 always holds two adjacent compass pixels (ring 0, 4, 8, 12), so m > t needs two
// adjacent compass pixels both darker or both brighter than the centre by more than t.  `pre2` = t + 256 in both halves.
template <int TP> __device__ __forceinline__ bool fastPretest(unsigned p, unsigned pre2) {
    constexpr int P = TP;
    const unsigned base = fastBase((int)ldsU8<0>(p));
    const unsigned e0 = fastE(ldsU8<3 * P>(p), base), e4 = fastE(ldsU8<3>(p), base);
    const unsigned e8 = fastE(ldsU8<-3 * P>(p), base), e12 = fastE(ldsU8<-3>(p), base);
    const unsigned pm = __vimax3_u16x2(__vminu2(e0, e4), __vminu2(e4, e8), __vmaxu2(__vminu2(e8, e12), __vminu2(e12, e0)));
    return __vmaxu2(pm, pre2) != pre2;
}

// Corner measure m (max over the 16 arcs of 9 contiguous ring pixels of min(v-p) resp. min(p-v)),
// OpenCV fast_score.cpp cornerScore<16> without the threshold clamp.
//
// Arc network (36 packed min/max instead of 16 arcs x 8): arcs k and k+1 share the 8 ring pixels k+1..k+8, so
//   max(arc_k, arc_k+1) = min(e_k+1 .. e_k+8, max(e_k, e_k+9))        (lattice identity)
// and the 8-pixel cores are built from pair minima lo2_j (j odd) and their pair minima lo4_j.
template <int TP> __device__ __forceinline__ int fastMeasure(unsigned p) {
    constexpr int P = TP;
    const unsigned base = fastBase((int)ldsU8<0>(p));
    unsigned e[16];
    e[0] = fastE(ldsU8<3 * P>(p), base); e[4] = fastE(ldsU8<3>(p), base);
    e[8] = fastE(ldsU8<-3 * P>(p), base); e[12] = fastE(ldsU8<-3>(p), base);
    e[1] = fastE(ldsU8<3 * P + 1>(p), base);   e[2] = fastE(ldsU8<2 * P + 2>(p), base);   e[3] = fastE(ldsU8<P + 3>(p), base);
    e[5] = fastE(ldsU8<-P + 3>(p), base);      e[6] = fastE(ldsU8<-2 * P + 2>(p), base);  e[7] = fastE(ldsU8<-3 * P + 1>(p), base);
    e[9] = fastE(ldsU8<-3 * P - 1>(p), base);  e[10] = fastE(ldsU8<-2 * P - 2>(p), base); e[11] = fastE(ldsU8<-P - 3>(p), base);
    e[13] = fastE(ldsU8<P - 3>(p), base);      e[14] = fastE(ldsU8<2 * P - 2>(p), base);  e[15] = fastE(ldsU8<3 * P - 1>(p), base);
    unsigned lo2[8], lo4[8], pr[8];   // index j <-> ring position 2j+1
#pragma unroll
    for (int j = 0; j < 8; j++) lo2[j] = __vminu2(e[2 * j + 1], e[(2 * j + 2) & 15]);
#pragma unroll
    for (int j = 0; j < 8; j++) lo4[j] = __vminu2(lo2[j], lo2[(j + 1) & 7]);
#pragma unroll
    for (int k = 0; k < 8; k++)      // arcs 2k and 2k+1: core = ring 2k+1 .. 2k+8 = lo4[k] and lo4[k+2]
        pr[k] = __vimin3_u16x2(lo4[k], lo4[(k + 2) & 7], __vmaxu2(e[2 * k], e[(2 * k + 9) & 15]));
    unsigned r = __vimax3_u16x2(__vimax3_u16x2(pr[0], pr[1], pr[2]), __vimax3_u16x2(pr[3], pr[4], pr[5]), __vmaxu2(pr[6], pr[7]));
    return (int)max(r & 0xffffu, r >> 16) - 256;
}

// NMS of cv::FAST is threshold-independent on m: a corner at threshold t (m > t) survives iff its
// stored score m-1 exceeds the stored scores of its 8 neighbours, and a neighbour's stored score is
// n-1 if n > t and 0 otherwise; since m > t >= 1 this is exactly "every neighbour has n < m".  The
// score tile holds m for m > t (t = the threshold in force) and 0 elsewhere (and outside the detection area).
template <bool PDL, int TP> __global__ void __launch_bounds__(kFastThreads, 32) k_fast(const __grid_constant__ FrameGeom G, const DevBuffers B) {
    if (PDL) pdlEnter(B);
    // dynamic shared memory, sized by the geometry's largest cell (FrameGeom::fastRegionA/fastScoreBytes/fastWarpCap):
    // [image tile][score tile][per-warp corner lists]
    extern __shared__ __align__(16) uint8_t fastSmem[];
    uint8_t *tile = fastSmem;
    uint8_t *score = fastSmem + G.fastRegionA;
    uint16_t *corners = reinterpret_cast<uint16_t *>(score + G.fastScoreBytes);   // per warp: tile offsets of pretest passers, then of the corners
    const int kFastWarpCap = G.fastWarpCap;
    const int f = blockIdx.y, tid = threadIdx.x;
    // cell window (src/ORBextractor.cc:749-762) from the host-built cell table; detection area = window minus 3 px
    const int2 ce = __ldg(B.tabs + G.cellTab + blockIdx.x);
    const int level = ce.x >> 24, iniX = ce.x & 0xfff, iniY = (ce.x >> 12) & 0xfff;
    const int dw = ce.y & 127, dh = (ce.y >> 7) & 127, cell = (int)((unsigned)ce.y >> 14);
    const int tw = dw + 6, th = dh + 6;              // window size
    const LevelGeom &L = G.lv[level];
    const int npx = dw * dh;
    const uint8_t *img = levelPtr(G, B, f, level) + (long long)iniY * L.pitch;
    // ---- tile load: aligned 32-bit words when the level allows it; tile column c <-> image x = xa + c
    const bool aligned = ((L.pitch & 3) == 0) && ((reinterpret_cast<uintptr_t>(img) & 3) == 0);
    const int xa = aligned ? (iniX & ~3) : iniX, ox = iniX - xa;
    const int lane = tid & 31, wid = tid >> 5;
    if (aligned) {
        // lane = word of a row (two rows per warp step when a row has <= 16 words): no index arithmetic in the loop,
        // and the unrolled loop keeps every load of the thread in flight at once
        const int nw = (ox + tw + 3) >> 2;           // words per row (<= 19)
        const bool two = nw <= 16;
        const int wq = two ? (lane & 15) : lane, r0 = two ? 2 * wid + (lane >> 4) : wid, rstep = two ? 2 * kFastWarps : kFastWarps;
        if (wq < nw) {
            const uint8_t *src = img + (long long)r0 * L.pitch + xa + 4 * wq;
            const long long sstep = (long long)rstep * L.pitch;
            uint8_t *dst = tile + r0 * TP + 4 * wq;
#pragma unroll 10
            for (int y = r0; y < th; y += rstep, src += sstep, dst += rstep * TP)
                *reinterpret_cast<uint32_t *>(dst) = __ldg(reinterpret_cast<const uint32_t *>(src));
        }
    } else {
        for (int y = wid; y < th; y += kFastThreads / 32) {
            const uint8_t *row = img + (long long)y * L.pitch + xa;
            for (int x = lane; x < tw; x += 32) tile[y * TP + x] = row[x];
        }
    }
    // score tile: all zero; phase 1 writes m only where m exceeds the threshold in force
    for (int i = tid; i < ((dh + 2) * TP + 15) / 16; i += kFastThreads) reinterpret_cast<uint4 *>(score)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    const int minTh = G.minTh, iniTh = G.iniTh;
    // ---- phase 1 runs at iniThFAST first and, only if no corner of the cell survives NMS, again at minThFAST
    // (src/ORBextractor.cc:764-770).  It is split in two so that the 36-op arc network only runs on full warps:
    //  1a  every detection pixel takes the compass pretest (5 loads, 7 packed min/max); the offsets of the pixels that
    //      pass go to the warp's own list (fill count in a register: no shared atomics, no shuffles).  On the textured
    //      frames 37 % pass at threshold 20 while 12 % are corners, so a warp that ran the full measure whenever one
    //      of its 32 lanes passed would never skip it;
    //  1b  the warp walks its list 32 entries at a time: full measure m, score tile entry and (in place, the write
    //      index never passes the read index) the list of the corners m > t.
    // Pixel i = y * dw + x advances by kFastThreads per iteration; y = i / dw comes from a 24-bit reciprocal (IMAD.HI on
    // i << 8, exact for i < 4096), so the address update runs on the FMA pipe and leaves the ALU pipe to the min/max ops.
    const unsigned ltMask = (1u << lane) - 1u;
    // 2^24 / dw rounded up by 1..11 (the float reciprocal is good to 2^-22): the quotient is exact while
    // i * (magic * dw - 2^24) < 2^24, i.e. up to an excess of 64 for dw <= 64
    const unsigned magic = (unsigned)(16777216.0f * __fdividef(1.0f, (float)dw)) + 6u;
    const int wrapOff = TP - dw;
    const unsigned centre = smemAddr(tile + 3 * TP + 3 + ox);
    const unsigned sc0a = smemAddr(score + TP + 1);
    const unsigned cornersAt = smemAddr(corners + wid * kFastWarpCap);
    uint2 *cand = B.cand + (long long)f * G.candPerFrame + L.candOff;
    int *counter = B.candCount + f * kMaxLevels + level;
    const int candCap = L.candCap;
    int t = iniTh;
    for (;;) {
        const unsigned pre2 = ((unsigned)t + 256u) * 0x10001u;
        int n0 = 0;
        for (int i = tid; i - lane < npx; i += kFastThreads) {
            const int ic = min(i, npx - 1);   // clamped, not branched: the tail lanes redo the last pixel and drop the result
            const int off = (int)__umulhi((unsigned)ic << 8, magic) * wrapOff + ic;
            const bool pass = fastPretest<TP>(centre + off, pre2) && i < npx;
            const unsigned mask = __ballot_sync(0xffffffffu, pass);
            if (pass) stsU16(cornersAt + 2 * (n0 + __popc(mask & ltMask)), (unsigned)off);
            n0 += __popc(mask);
        }
        __syncwarp();
        int n1 = 0;
        for (int k0 = 0; k0 < n0; k0 += 32) {
            const int k = k0 + lane;
            unsigned off = 0;
            int m = -1;
            if (k < n0) {
                off = ldsU16(cornersAt + 2 * k);
                m = fastMeasure<TP>(centre + off);
            }
            const bool corner = m > t;
            if (corner) stsU8(sc0a + off, (unsigned)m);
            const unsigned mask = __ballot_sync(0xffffffffu, corner);   // every lane has consumed its list entry by now
            if (corner) stsU16(cornersAt + 2 * (n1 + __popc(mask & ltMask)), off);
            n1 += __popc(mask);
        }
        __syncthreads();
        // ---- phase 2: strict local maxima of m among the warp's corners (non-corners hold 0 in the score tile) are the
        // keypoints of the cell.  A warp that finds one knows the pass is final (the cell has a survivor), so it appends its
        // maxima to the level's candidate array right away, one atomicAdd per batch; a cell where no warp finds any repeats
        // the pass at minThFAST.
        int any = 0;
        for (int k0 = 0; k0 < n1; k0 += 32) {
            const int k = k0 + lane;
            bool ismax = false;
            unsigned o = 0, m = 0;
            if (k < n1) {
                constexpr int P = TP;
                o = ldsU16(cornersAt + 2 * k);
                const unsigned sc = sc0a + o;
                m = ldsU8<0>(sc);
                const unsigned nb = max(max(max(ldsU8<-1>(sc), ldsU8<1>(sc)), max(ldsU8<-P - 1>(sc), ldsU8<-P>(sc))),
                                        max(max(ldsU8<-P + 1>(sc), ldsU8<P - 1>(sc)), max(ldsU8<P>(sc), ldsU8<P + 1>(sc))));
                ismax = nb < m;
            }
            const unsigned mask = __ballot_sync(0xffffffffu, ismax);
            if (mask == 0) continue;
            any = 1;
            int base = 0;
            if (lane == 0) base = atomicAdd(counter, __popc(mask));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (ismax) {
                const int slot = base + __popc(mask & ltMask);
                if (slot < candCap) {
                    const unsigned y = (o * ((1u << 20) / TP + 1)) >> 20, x = o - y * TP;   // exact for o < 2^13
                    const uint32_t ax = iniX + 3 + x, ay = iniY + 3 + y;  // absolute level coordinates
                    cand[slot] = make_uint2(ax | (ay << 12) | ((uint32_t)(m - 1) << 24), ((uint32_t)cell << 12) | (y << 6) | x);
                } else {
                    atomicOr(B.status, 1);
                }
            }
        }
        if (__syncthreads_or(any) || t == minTh) break;
        t = minTh;
    }
}

int launchFast(const FrameGeom &G, const DevBuffers &B, int nFrames, cudaStream_t s) {
    if (G.totalCells == 0) return 0;
    const size_t smem = (size_t)G.fastRegionA + G.fastScoreBytes + (size_t)(kFastThreads / 32) * G.fastWarpCap * sizeof(uint16_t);
    const dim3 grid(G.totalCells, nFrames);
    if (G.fastPitch == kFastTilePitchNarrow) {
        if (B.pdl) launchChain(true, k_fast<true, kFastTilePitchNarrow>, grid, dim3(kFastThreads), smem, s, G, B);
        else k_fast<false, kFastTilePitchNarrow><<<grid, kFastThreads, smem, s>>>(G, B);
    } else {
        if (B.pdl) launchChain(true, k_fast<true, kFastTilePitch>, grid, dim3(kFastThreads), smem, s, G, B);
        else k_fast<false, kFastTilePitch><<<grid, kFastThreads, smem, s>>>(G, B);
    }
    return 1;
}

// =================================================================================================
// K3 — DistributeOctTree (src/ORBextractor.cc:533-723) + DivideNode (:479-531), one CTA per
// (frame, level), reproducing the reference's node list ORDER and split order exactly.
//
// The reference is a serial std::list algorithm; what it computes can be stated without the list:
//  * a pass splits a set of parent nodes in a processing order r = 0..P-1; every non-empty child is
//    push_front'ed in n1..n4 order, so after the pass the list reads: children of the LAST processed
//    parent (n4..n1), ..., children of the first, then the untouched nodes in their old order;
//  * phase 1 (:591-643) processes every multi-point node in list order; phase 2 (:649-699) takes the
//    multi-point children of the previous pass, sorts them by (size, pointer) and splits from the
//    largest down, stopping as soon as the list holds N nodes.  Pointer order is canonicalised to
//    creation order (SURVEY.md §7 hard part 1), i.e. the push order of the previous pass (`seq`);
//  * the final pick per node (:703-720) is the max response, first in key order on ties; key order
//    inside a node is the original vToDistributeKeys order, carried here by the candidates' order key.
// So points never move: each point carries the index of its node, and a pass is two sweeps over the
// points (count children, relabel) plus block-wide scans over the node list.
#ifndef YGZ_QT_SINGLE
#define YGZ_QT_SINGLE 1024
#endif
constexpr int kQtThreads = YGZ_QT_SINGLE;        // CTA size for small batches (one frame: latency; 1024: 41 us, 512: 44, 256: 56 on 640x480); batches run kQtThreadsBatch
constexpr int kQtThreadsBatch = 384;   // more CTAs per SM hide the barrier stalls of the passes (0.47 -> 0.42 ms / 512 frames at 384 threads; 256: 0.43, 192: 0.47, 448: 0.59)

struct QtSmem {
    short4 *bA, *bB;      // node bounds x0,x1,y0,y1 (x = UL.x, y = UR.x, z = UL.y, w = BL.y)
    int *cA, *cB;         // points per node
    int *gA, *gB;         // seq (creation order in the latest pass) if multi-point child of it, else -1
    int *procOrder;       // per list node: processing rank r or -1
    int *byOrder;         // rank -> list node
    int *childCnt;        // [4*r + q]
    int *childIdx;        // [4*r + q] -> new list index
    int *oldToNew;        // untouched node -> new list index
    int *scanR, *scanU;   // scan scratch
    int *misc;            // [0..31] warp sums, [32] nToExpand
};

__device__ __forceinline__ int qtQuadrant(uint32_t xyr, const short4 b) {
    // DivideNode (:481-482, 512-523): halfX = ceil((UR.x-UL.x)/2), child by pt.x < UL.x+halfX etc.
    const int x = (int)(xyr & 0xfff) - kMinBorder, y = (int)((xyr >> 12) & 0xfff) - kMinBorder;
    const int mx = b.x + ((b.y - b.x + 1) >> 1), my = b.z + ((b.w - b.z + 1) >> 1);
    return (x < mx ? 0 : 1) + (y < my ? 0 : 2);
}

// In-place exclusive scan of a[0..n) by the whole CTA; returns the total.  `ws` = 33 ints scratch.
template <int QT> __device__ int qtScan(int *a, int n, int *ws) {
    const int T = QT, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int chunk = (n + T - 1) / T;
    const int b = min(tid * chunk, n), e = min(b + chunk, n);
    int sum = 0;
    for (int i = b; i < e; i++) sum += a[i];
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) ws[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int w = lane < T / 32 ? ws[lane] : 0;
        int winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += v;
        }
        if (lane < T / 32) ws[lane] = winc - w;
        if (lane == 31) ws[32] = winc;
    }
    __syncthreads();
    int run = ws[wid] + inc - sum;
    for (int i = b; i < e; i++) { const int v = a[i]; a[i] = run; run += v; }
    const int total = ws[32];
    __syncthreads();
    return total;
}

// Sweep 1 of a pass: per processed parent (rank r), count the points falling in each quadrant.
template <int QT> __device__ void qtCountChildren(const QtSmem &S, const short4 *bnd, int P, const uint2 *cand, const uint16_t *nodeOf, int np) {
    for (int i = threadIdx.x; i < 4 * P; i += QT) S.childCnt[i] = 0;
    __syncthreads();
    for (int p = threadIdx.x; p < np; p += QT) {
        const int nd = nodeOf[p];
        const int r = S.procOrder[nd];
        if (r >= 0) atomicAdd(&S.childCnt[4 * r + qtQuadrant(cand[p].x, bnd[nd])], 1);
    }
    __syncthreads();
}

// Rest of a pass: builds the new list (into b2/c2/g2) from the old one (b1/c1) given procOrder /
// byOrder / childCnt for P processed parents, relabels the points.  Returns the new list size and
// the number of multi-point children through *nToExpand.
template <int QT> __device__ int qtCommit(const QtSmem &S, const short4 *b1, const int *c1, short4 *b2, int *c2, int *g2, int n, int P,
                        const uint2 *cand, uint16_t *nodeOf, int np, int *nToExpand) {
    const int tid = threadIdx.x;
    for (int r = tid; r < P; r += QT) {
        const int *c = &S.childCnt[4 * r];
        S.scanR[r] = (c[0] > 0) + (c[1] > 0) + (c[2] > 0) + (c[3] > 0);
    }
    for (int i = tid; i < n; i += QT) S.scanU[i] = S.procOrder[i] < 0;
    if (tid == 0) S.misc[33] = 0;
    __syncthreads();
    const int totalChildren = qtScan<QT>(S.scanR, P, S.misc);
    const int untouched = qtScan<QT>(S.scanU, n, S.misc);
    for (int r = tid; r < P; r += QT) {
        const short4 pb = b1[S.byOrder[r]];
        const int mx = pb.x + ((pb.y - pb.x + 1) >> 1), my = pb.z + ((pb.w - pb.z + 1) >> 1);
        int seq = S.scanR[r], multi = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const int cnt = S.childCnt[4 * r + q];
            if (cnt > 0) {
                const int pos = totalChildren - 1 - seq;
                short4 cb;
                cb.x = (q & 1) ? mx : pb.x; cb.y = (q & 1) ? pb.y : mx;
                cb.z = (q & 2) ? my : pb.z; cb.w = (q & 2) ? pb.w : my;
                b2[pos] = cb; c2[pos] = cnt; g2[pos] = cnt > 1 ? seq : -1;
                S.childIdx[4 * r + q] = pos;
                multi += cnt > 1;
                seq++;
            }
        }
        if (multi) atomicAdd(&S.misc[33], multi);
    }
    for (int i = tid; i < n; i += QT) {
        if (S.procOrder[i] < 0) {
            const int pos = totalChildren + S.scanU[i];
            b2[pos] = b1[i]; c2[pos] = c1[i]; g2[pos] = -1;
            S.oldToNew[i] = pos;
        }
    }
    __syncthreads();
    for (int p = tid; p < np; p += QT) {
        const int nd = nodeOf[p];
        const int r = S.procOrder[nd];
        nodeOf[p] = (uint16_t)(r >= 0 ? S.childIdx[4 * r + qtQuadrant(cand[p].x, b1[nd])] : S.oldToNew[nd]);
    }
    *nToExpand = S.misc[33];
    __syncthreads();
    return totalChildren + untouched;
}

template <bool PDL, int QT> __global__ void __launch_bounds__(QT) k_quadtree(const __grid_constant__ FrameGeom G, const DevBuffers B) {
    if (PDL) pdlEnter(B);
    extern __shared__ __align__(16) unsigned char qtRaw[];
    const int level = blockIdx.x, f = blockIdx.y, tid = threadIdx.x;
    const LevelGeom &L = G.lv[level];
    int *keptCount = B.keptCount + f * kMaxLevels + level;
    const int np = L.nCols == 0 ? 0 : min(B.candCount[f * kMaxLevels + level], L.candCap);
    if (np == 0) { if (tid == 0) *keptCount = 0; return; }
    const int C = G.maxListCap;
    QtSmem S;
    {
        unsigned char *p = qtRaw;
        S.bA = (short4 *)p; p += sizeof(short4) * C;
        S.bB = (short4 *)p; p += sizeof(short4) * C;
        S.cA = (int *)p; p += 4 * C;   S.cB = (int *)p; p += 4 * C;
        S.gA = (int *)p; p += 4 * C;   S.gB = (int *)p; p += 4 * C;
        S.procOrder = (int *)p; p += 4 * C;   S.byOrder = (int *)p; p += 4 * C;
        S.childCnt = (int *)p; p += 16 * C;   S.childIdx = (int *)p; p += 16 * C;
        S.oldToNew = (int *)p; p += 4 * C;
        S.scanR = (int *)p; p += 4 * C;   S.scanU = (int *)p; p += 4 * C;
        S.misc = (int *)p;
    }
    const uint2 *cand = B.cand + (long long)f * G.candPerFrame + L.candOff;
    uint16_t *nodeOf = B.nodeOf + (long long)f * G.candPerFrame + L.candOff;
    const int N = L.nTarget;
    short4 *b1 = S.bA, *b2 = S.bB;
    int *c1 = S.cA, *c2 = S.cB, *g1 = S.gA, *g2 = S.gB;

    // ---- roots (:537-574): nIni roots of width hX over [0, maxX-minX) x [0, maxY-minY]; empty ones dropped
    const int nIni = L.nIni;
    const float hX = L.hX;
    for (int i = tid; i < nIni; i += QT) S.childCnt[i] = 0;
    __syncthreads();
    for (int p = tid; p < np; p += QT) {
        const int x = (int)(cand[p].x & 0xfff) - kMinBorder;
        const int root = (int)__fdiv_rn((float)x, hX);
        nodeOf[p] = (uint16_t)root;
        atomicAdd(&S.childCnt[root], 1);
    }
    __syncthreads();
    for (int i = tid; i < nIni; i += QT) S.scanU[i] = S.childCnt[i] > 0;
    __syncthreads();
    int n = qtScan<QT>(S.scanU, nIni, S.misc);
    for (int i = tid; i < nIni; i += QT) {
        const int cnt = S.childCnt[i];
        if (cnt > 0) {
            const int pos = S.scanU[i];
            b1[pos] = make_short4((short)(int)__fmul_rn(hX, (float)i), (short)(int)__fmul_rn(hX, (float)(i + 1)), 0,
                                  (short)(L.h - 2 * kMinBorder));
            c1[pos] = cnt; g1[pos] = -1;
        }
    }
    __syncthreads();
    for (int p = tid; p < np; p += QT) nodeOf[p] = (uint16_t)S.scanU[nodeOf[p]];
    __syncthreads();

    // ---- main loop (:576-700)
    bool finish = false;
    while (!finish) {
        const int prevSize = n;
        // phase-1 pass: every multi-point node, in list order
        for (int i = tid; i < n; i += QT) S.scanR[i] = c1[i] > 1;
        __syncthreads();
        int P = qtScan<QT>(S.scanR, n, S.misc);
        for (int i = tid; i < n; i += QT) {
            const bool e = c1[i] > 1;
            S.procOrder[i] = e ? S.scanR[i] : -1;
            if (e) S.byOrder[S.scanR[i]] = i;
        }
        __syncthreads();
        int nToExpand = 0;
        if (P > 0) {
            qtCountChildren<QT>(S, b1, P, cand, nodeOf, np);
            n = qtCommit<QT>(S, b1, c1, b2, c2, g2, n, P, cand, nodeOf, np, &nToExpand);
            { short4 *t = b1; b1 = b2; b2 = t; } { int *t = c1; c1 = c2; c2 = t; } { int *t = g1; g1 = g2; g2 = t; }
        }
        if (n >= N || n == prevSize) {
            finish = true;
        } else if (n + nToExpand * 3 > N) {
            // phase 2 (:649-699): latest multi-point children, largest first (later-created first among equals)
            while (!finish) {
                const int prevSize2 = n;
                // compact the generation: gen entry e -> list node, keyed by (size, seq)
                for (int i = tid; i < n; i += QT) S.scanR[i] = g1[i] >= 0;
                __syncthreads();
                const int F = qtScan<QT>(S.scanR, n, S.misc);
                for (int i = tid; i < n; i += QT) if (g1[i] >= 0) S.oldToNew[S.scanR[i]] = i;
                __syncthreads();
                // rank r = number of generation nodes that sort after this one (processed earlier)
                for (int i = tid; i < n; i += QT) S.procOrder[i] = -1;
                __syncthreads();
                for (int e = tid; e < F; e += QT) {
                    const int i = S.oldToNew[e];
                    const int sz = c1[i], sq = g1[i];
                    int r = 0;
                    for (int e2 = 0; e2 < F; e2++) {
                        const int j = S.oldToNew[e2];
                        const int sz2 = c1[j];
                        r += (sz2 > sz) || (sz2 == sz && g1[j] > sq);
                    }
                    S.procOrder[i] = r;
                    S.byOrder[r] = i;
                }
                __syncthreads();
                qtCountChildren<QT>(S, b1, F, cand, nodeOf, np);
                // the split loop breaks once the list holds N nodes: rank r is processed iff the list
                // size after ranks < r is still below N
                for (int r = tid; r < F; r += QT) {
                    const int *c = &S.childCnt[4 * r];
                    S.scanU[r] = (c[0] > 0) + (c[1] > 0) + (c[2] > 0) + (c[3] > 0) - 1;
                }
                __syncthreads();
                qtScan<QT>(S.scanU, F, S.misc);   // exclusive: gain of ranks < r
                if (tid == 0) S.misc[34] = 0;
                __syncthreads();
                int cntProc = 0;
                for (int r = tid; r < F; r += QT) cntProc += (n + S.scanU[r] < N);
                if (cntProc) atomicAdd(&S.misc[34], cntProc);
                __syncthreads();
                P = S.misc[34];   // processed ranks are a prefix 0..P-1 (gains are >= 0)
                for (int r = tid; r < F; r += QT) if (r >= P) S.procOrder[S.byOrder[r]] = -1;
                __syncthreads();
                n = qtCommit<QT>(S, b1, c1, b2, c2, g2, n, P, cand, nodeOf, np, &nToExpand);
                { short4 *t = b1; b1 = b2; b2 = t; } { int *t = c1; c1 = c2; c2 = t; } { int *t = g1; g1 = g2; g2 = t; }
                if (n >= N || n == prevSize2) finish = true;
            }
        }
    }

    // ---- best response per node, first in key order on ties (:703-720); output in list order
    unsigned long long *best = (unsigned long long *)S.childCnt;
    for (int i = tid; i < n; i += QT) best[i] = 0ull;
    __syncthreads();
    for (int p = tid; p < np; p += QT) {
        const uint2 c = cand[p];
        const unsigned long long key = ((unsigned long long)(c.x >> 24) << 32) | (unsigned long long)(0xffffffffu - c.y);
        atomicMax(&best[nodeOf[p]], key);
    }
    __syncthreads();
    uint32_t *kept = B.kept + (long long)f * G.keptCap + L.keptOff;
    for (int p = tid; p < np; p += QT) {
        const uint2 c = cand[p];
        const unsigned long long key = ((unsigned long long)(c.x >> 24) << 32) | (unsigned long long)(0xffffffffu - c.y);
        const int nd = nodeOf[p];
        if (best[nd] == key && nd < L.listCap) kept[nd] = c.x;
    }
    if (tid == 0) {
        *keptCount = min(n, L.listCap);
        if (n > L.listCap) atomicOr(B.status, 2);
    }
}

size_t quadtreeSmemBytes(int listCap) {
    const size_t C = (size_t)listCap;
    return sizeof(short4) * 2 * C + 4 * C * 4 + 4 * C * 2 + 16 * C * 2 + 4 * C * 3 + 40 * 4;
}

int launchQuadtree(const FrameGeom &G, const DevBuffers &B, int nFrames, cudaStream_t s) {
    const size_t smem = quadtreeSmemBytes(G.maxListCap);
    if (B.pdl) launchChain(true, k_quadtree<true, kQtThreads>, dim3(G.nlevels, nFrames), dim3(kQtThreads), smem, s, G, B);
    else if (nFrames <= 8) k_quadtree<false, kQtThreads><<<dim3(G.nlevels, nFrames), kQtThreads, smem, s>>>(G, B);
    else k_quadtree<false, kQtThreadsBatch><<<dim3(G.nlevels, nFrames), kQtThreadsBatch, smem, s>>>(G, B);
    return 1;
}

// =================================================================================================
// K5 — cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) on every level (src/ORBextractor.cc:1009-1010),
// OpenCV >= 3.4 fixed-point path: taps {18,34,48,56,48,34,18}/256, horizontal sum in 16 bits,
// vertical sum in 32 bits, one rounding (v + 32768) >> 16.  Shared-memory tile with a 3-px halo;
// one launch covers all levels of all frames (tiles enumerated level-major).
// Tile = kBlurTW x kBlurTH outputs per CTA (ygz_common.cuh); 256 threads = 64 column groups of 4 px x 4 row groups.
constexpr int kBlurThreads = 256;
constexpr int kBlurRawRows = kBlurTH + 6;

__device__ __forceinline__ int reflect101(int p, int len) {
    if (len == 1) return 0;
    while ((unsigned)p >= (unsigned)len) p = p < 0 ? -p : 2 * len - 2 - p;
    return p;
}

constexpr int kBlurRawPitch = kBlurTW + 32;     // bytes; raw column c <-> image x = x0 - 16 + c (16-byte aligned chunks)


constexpr int kBlurPairs = kBlurRawRows / 2;    // 19 raw-row pairs (rows 2m, 2m+1)
static_assert(kBlurRawRows % 2 == 0 && kBlurTH % 8 == 0 && kBlurTW == 256, "k_blur thread mapping");

// byte 2 of four 24-bit accumulators -> one word of four output bytes
__device__ __forceinline__ uint32_t packByte2(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3) {
    return __byte_perm(__byte_perm(a0, a1, 0x0062), __byte_perm(a2, a3, 0x0062), 0x5410);
}

template <bool PDL> __global__ void __launch_bounds__(kBlurThreads, 6) k_blur(const __grid_constant__ FrameGeom G, const DevBuffers B) {
    if (PDL) pdlEnter(B);
    __shared__ __align__(16) uint8_t raw[kBlurRawRows * kBlurRawPitch];
    // horizontal sums, stored as ROW PAIRS: hs[m][g] = four columns (4g .. 4g+3) of h[row 2m] | h[row 2m+1] << 16 — the
    // operand form of DP2A in the vertical pass, so that pass needs no re-packing
    __shared__ __align__(16) uint4 hs[kBlurPairs * (kBlurTW / 4)];
    const int f = blockIdx.y, tid = threadIdx.x;
    const int2 te = __ldg(B.tabs + G.tileTab + blockIdx.x);            // host-built tile table
    const int level = te.x >> 24, x0 = te.x & 0xfff, y0 = (te.x >> 12) & 0xfff;
    const LevelGeom &L = G.lv[level];
    const uint8_t *img = levelPtr(G, B, f, level);
    // ---- stage the raw tile (rows y0-3 .. y0+34, columns x0-3 .. x0+258 needed; REFLECT_101 about the true image
    // edge).  16-byte chunks that lie inside the image row go global -> shared with cp.async (no registers, all in
    // flight at once); the columns left over at the image's left/right edge (<= 3 + 18 per row) are reflected
    // byte-wise.  Tiny or 16-byte-unaligned levels take the generic byte path.
    constexpr int kChunks = kBlurRawPitch / 16;   // 18
    const bool aligned = ((L.pitch & 15) == 0) && ((reinterpret_cast<uintptr_t>(img) & 15) == 0);
    if (aligned && L.w >= 8 && L.h >= 8) {
        const int hm = 2 * (L.h - 1), wm = 2 * (L.w - 1);
        // thread = (chunk column c, row r0 + 14 i): 252 threads cover 14 rows per step, no division in the loop
        constexpr int kRowsPerStep = kBlurThreads / kChunks;   // 14
        const int r0 = tid / kChunks, c = tid - r0 * kChunks;
        const int gx = x0 - 16 + 16 * c;
        if (r0 < kRowsPerStep && gx >= 0 && gx + 16 <= L.w) {
            const uint8_t *src = img + gx;
            uint8_t *dst = &raw[r0 * kBlurRawPitch + 16 * c];
#pragma unroll
            for (int i = 0; i < (kBlurRawRows + kRowsPerStep - 1) / kRowsPerStep; i++) {
                const int r = r0 + kRowsPerStep * i;
                if (r < kBlurRawRows) {
                    int sy = abs(y0 + r - 3);
                    sy = max(min(sy, hm - sy), 0);                   // one reflection: the halo (3) is shorter than the image
                    cpAsync16(dst + i * kRowsPerStep * kBlurRawPitch, src + (long long)sy * L.pitch);
                }
            }
        }
        // columns the chunks did not bring: [-3, 0) of the image, and [w & ~15, w + 3) clipped to this tile
        const int nLeft = x0 == 0 ? 3 : 0;
        const int rb = max(L.w & ~15, x0 - 3), re = min(L.w + 3, x0 + kBlurTW + 3);
        const int nb = nLeft + max(re - rb, 0);
        if (nb > 0) {
            const float rnb = __fdividef(1.0f, (float)nb);
            for (int i = tid; i < kBlurRawRows * nb; i += kBlurThreads) {
                const int r = (int)(((float)i + 0.5f) * rnb), k = i - r * nb;
                const int col = k < nLeft ? k - 3 : rb + (k - nLeft);
                int sy = abs(y0 + r - 3);
                sy = max(min(sy, hm - sy), 0);
                int sx = abs(col);
                sx = max(min(sx, wm - sx), 0);
                raw[r * kBlurRawPitch + col - x0 + 16] = img[(long long)sy * L.pitch + sx];
            }
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
    } else {
        constexpr int kCols = kBlurTW + 8;        // columns x0-4 .. x0+259
        for (int i = tid; i < kBlurRawRows * kCols; i += kBlurThreads) {
            const int r = i / kCols, c = i - r * kCols;
            const int x = x0 - 4 + c;
            uint8_t v = 0;
            if (x >= -3 && x <= L.w + 2) v = img[(long long)reflect101(y0 + r - 3, L.h) * L.pitch + reflect101(x, L.w)];
            raw[r * kBlurRawPitch + 12 + c] = v;
        }
    }
    __syncthreads();
    // ---- horizontal pass: thread = 4 columns (x .. x+3) of one raw-row PAIR.  Per row the words w0 = [x-4, x),
    // w1 = [x, x+4), w2 = [x+4, x+8); output k is the dot product of bytes (1+k .. 7+k) of that window with
    // {18,34,48,56,48,34,18}: two byte-aligned windows (funnel shifts) and two DP4A.  The two rows of the pair are packed
    // per column (low half = even raw row).  Warps whose 128 columns lie right of the image skip both passes.
    const int tx = tid & 63, tg = tid >> 6;
    const int x = x0 + 4 * tx;
    const bool live = x0 + 4 * (tx & ~31) < L.w;        // warp-uniform
    constexpr uint32_t kW0 = 18u | (34u << 8) | (48u << 16) | (56u << 24), kW1 = 48u | (34u << 8) | (18u << 16);
    if (live) {
        const uint32_t *w = reinterpret_cast<const uint32_t *>(&raw[2 * tg * kBlurRawPitch + 12 + 4 * tx]);
        uint4 *hd = hs + tg * (kBlurTW / 4) + tx;
#pragma unroll
        for (int i = 0; i < (kBlurPairs + 3) / 4; i++) {
            if (tg + 4 * i < kBlurPairs) {
                const uint32_t *wa = w + i * (8 * kBlurRawPitch / 4), *wb = wa + kBlurRawPitch / 4;
                const uint32_t a0 = wa[0], a1 = wa[1], a2 = wa[2], b0 = wb[0], b1 = wb[1], b2 = wb[2];
                uint32_t o[4];
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    const uint32_t ha = __dp4a(__funnelshift_r(a1, a2, 8 * (k + 1)), kW1, __dp4a(__funnelshift_r(a0, a1, 8 * (k + 1)), kW0, 0u));
                    const uint32_t hb = __dp4a(__funnelshift_r(b1, b2, 8 * (k + 1)), kW1, __dp4a(__funnelshift_r(b0, b1, 8 * (k + 1)), kW0, 0u));
                    o[k] = __byte_perm(ha, hb, 0x5410);
                }
                o[3] = __byte_perm(__dp4a(a2, kW1, __dp4a(a1, kW0, 0u)), __dp4a(b2, kW1, __dp4a(b1, kW0, 0u)), 0x5410);
                hd[i * 4 * (kBlurTW / 4)] = make_uint4(o[0], o[1], o[2], o[3]);
            }
        }
    }
    __syncthreads();
    // ---- vertical pass: thread = 4 columns x 8 output rows (k = 8 tg .. 8 tg + 7), from the 7 row pairs Q[4 tg .. 4 tg + 6].
    // Output row k covers raw rows k .. k+6.  Even k = 2m: pairs m, m+1, m+2 whole and the low half of m+3; odd k = 2m+1:
    // the high half of pair m, then m+1, m+2, m+3 whole with the taps shifted by one.  Four DP2A per output either way, the
    // rounding constant rides in the accumulator, the result is byte 2 of the 24-bit sum.
    if (!live || x >= L.w) return;
    uint8_t *dst = B.blur + (long long)f * B.blurSlab + B.blurOff[level] + (long long)(y0 + 8 * tg) * B.blurPitch[level] + x;
    const int dp = B.blurPitch[level];
    const int nrows = L.h - (y0 + 8 * tg);            // rows of this thread that exist (may be <= 0)
    uint4 Q[7];
#pragma unroll
    for (int j = 0; j < 7; j++) Q[j] = hs[(4 * tg + j) * (kBlurTW / 4) + tx];
    constexpr uint32_t kE0 = 18u | (34u << 8), kE1 = 48u | (56u << 8), kE2 = 48u | (34u << 8), kE3 = 18u;
    constexpr uint32_t kO0 = 18u << 8, kO1 = 34u | (48u << 8), kO2 = 56u | (48u << 8), kO3 = 34u | (18u << 8);
#define YGZ_BLUR_COL(F, W0, W1, W2, W3) \
    __dp2a_lo(Q[mm + 3].F, W3, __dp2a_lo(Q[mm + 2].F, W2, __dp2a_lo(Q[mm + 1].F, W1, __dp2a_lo(Q[mm].F, W0, 32768u))))
#pragma unroll
    for (int mm = 0; mm < 4; mm++) {
        const uint32_t ev = packByte2(YGZ_BLUR_COL(x, kE0, kE1, kE2, kE3), YGZ_BLUR_COL(y, kE0, kE1, kE2, kE3),
                                      YGZ_BLUR_COL(z, kE0, kE1, kE2, kE3), YGZ_BLUR_COL(w, kE0, kE1, kE2, kE3));
        const uint32_t od = packByte2(YGZ_BLUR_COL(x, kO0, kO1, kO2, kO3), YGZ_BLUR_COL(y, kO0, kO1, kO2, kO3),
                                      YGZ_BLUR_COL(z, kO0, kO1, kO2, kO3), YGZ_BLUR_COL(w, kO0, kO1, kO2, kO3));
        if (2 * mm < nrows) *reinterpret_cast<uint32_t *>(dst + (2 * mm) * dp) = ev;
        if (2 * mm + 1 < nrows) *reinterpret_cast<uint32_t *>(dst + (2 * mm + 1) * dp) = od;
    }
#undef YGZ_BLUR_COL
}

int launchBlur(const FrameGeom &G, const DevBuffers &B, int nFrames, cudaStream_t s) {
    if (G.totalTiles == 0) return 0;
    if (B.pdl) launchChain(true, k_blur<true>, dim3(G.totalTiles, nFrames), dim3(kBlurThreads), 0, s, G, B);
    else k_blur<false><<<dim3(G.totalTiles, nFrames), kBlurThreads, 0, s>>>(G, B);
    return 1;
}

// =================================================================================================
// K4 + K6 — IC_Angle (src/ORBextractor.cc:77-101) on the unblurred level, then computeOrbDescriptor
// (:105-149) on the blurred level, one warp per kept keypoint; also the final assembly (:1019-1026):
// level-major concatenation, pt *= mvScaleFactor[level] for level > 0.
//
// cosf/sinf: the reference calls cos(float)/sin(float) under `using namespace std` (:67,108-109),
// i.e. glibc's float routines.  devCosf/devSinf restate glibc 2.39's s_cosf.c/s_sinf.c (double
// polynomial with fused multiply-adds, x86-64 FMA variant) op for op — IEEE double fma on the GPU
// gives the same bits.  cv::fastAtan2 (OpenCV mathfuncs_core atan_f32) is restated in float with
// explicit round-to-nearest mul/add so no contraction happens.
__device__ __forceinline__ float sincosPolyDev(double x, double x2, int n, bool neg) {
    const double C0 = 1.0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10,
                 C4 = 0x1.99343027bf8c3p-16;
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    if ((n & 1) == 0) {
        const double x3 = __dmul_rn(x, x2);
        const double s1 = __fma_rn(x2, S3, S2);
        const double x7 = __dmul_rn(x3, x2);
        const double s = __fma_rn(x3, S1, x);
        return (float)__fma_rn(x7, s1, s);
    }
    const double sg = neg ? -1.0 : 1.0;
    const double x4 = __dmul_rn(x2, x2);
    const double c2 = __fma_rn(x2, sg * C4, sg * C3);
    const double c1 = __fma_rn(x2, sg * C1, sg * C0);
    const double x6 = __dmul_rn(x4, x2);
    const double c = __fma_rn(x4, sg * C2, c1);
    return (float)__fma_rn(x6, c2, c);
}
__device__ __forceinline__ void devSinCosf(float y, float *sn, float *cs) {
    const uint32_t top = (__float_as_uint(y) >> 20) & 0x7ff;
    double x = (double)y;
    if (top < 0x3f4) {  // |y| < pi/4
        const double x2 = __dmul_rn(x, x);
        if (top < 0x398) { *sn = y; *cs = 1.0f; return; }
        *sn = sincosPolyDev(x, x2, 0, false);
        *cs = sincosPolyDev(x, x2, 1, false);
        return;
    }
    const double r = __dmul_rn(x, 0x1.45f306dc9c883p+23);
    const int n = (__double2int_rz(r) + 0x800000) >> 24;
    x = __fma_rn(-(double)n, 0x1.921fb54442d18p+0, x);
    const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;
    const double xs = __dmul_rn(x, sgn), x2 = __dmul_rn(x, x);
    *sn = sincosPolyDev(xs, x2, n, (n & 2) != 0);
    *cs = sincosPolyDev(xs, x2, n ^ 1, (n & 2) != 0);
}

__device__ __forceinline__ float devFastAtan2(float y, float x) {
    const float k = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * k, p3 = -0.3258083974640975f * k, p5 = 0.1555786518463281f * k,
                p7 = -0.04432655554792128f * k;
    const float eps = 2.2204460492503131e-16f;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

__device__ __align__(16) signed char gPattern[1024];  // filled from kOrbPattern by configureKernels()

// IC_Angle (:79-109) by one warp: lane <-> column u = lane-15; the rows of the circular patch that hold column u are
// |v| <= vmaxOfU[|u|], taken in +-v pairs like the reference (integer sums: order-free).  c0 = patch centre.
__device__ __forceinline__ float icAngleWarp(const FrameGeom &G, const uint8_t *c0, int pitch, int lane) {
    int m10 = 0, m01 = 0;
    const int u = lane - kHalfPatch;
    if (lane < 31) {
        const int vmax = G.vmaxOfU[abs(u)];
        const uint8_t *up = c0 + u, *dn = up;
        int sum = *up;
#pragma unroll 4
        for (int v = 1; v <= vmax; v++) {
            up -= pitch; dn += pitch;
            const int a = *dn, b = *up;
            sum += a + b;
            m01 += v * (a - b);
        }
        m10 = u * sum;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    return devFastAtan2((float)m01, (float)m10);
}

// computeOrbDescriptor (:105-149) by one warp: lane <-> byte, 8 tests each.  cb = patch centre on the blurred level.
__device__ __forceinline__ int orbByteWarp(const uint8_t *cb, int bp, float angle, int lane) {
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    float a, b;
    devSinCosf(__fmul_rn(angle, factorPI), &b, &a);
    const char4 *pat = reinterpret_cast<const char4 *>(gPattern) + lane * 8;  // 8 tests (x0,y0,x1,y1) per byte
    int val = 0;
#pragma unroll
    for (int t = 0; t < 8; t++) {
        const char4 p = pat[t];
        // GET_VALUE(idx) = center[cvRound(x*b + y*a)*step + cvRound(x*a - y*b)]  (:114-116)
        const float x0 = p.x, y0 = p.y, x1 = p.z, y1 = p.w;
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a)));
        const int q0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a)));
        const int q1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)));
        const int t0 = cb[r0 * bp + q0], t1 = cb[r1 * bp + q1];
        val |= (t0 < t1) << t;
    }
    return val;
}

constexpr int kDescWarps = 4;

template <bool PDL> __global__ void __launch_bounds__(kDescWarps * 32) k_describe(const __grid_constant__ FrameGeom G, const DevBuffers B, int cap,
                                                              int32_t *counts, ygzorb_keypoint *kps, uint8_t *desc) {
    if (PDL) pdlEnter(B);
    const int f = blockIdx.y, lane = threadIdx.x & 31;
    const int j = blockIdx.x * kDescWarps + (threadIdx.x >> 5);
    // level-major offsets from the per-level kept counts
    const int *kc = B.keptCount + f * kMaxLevels;
    int level = -1, k = 0, total = 0;
    for (int l = 0; l < G.nlevels; l++) {
        const int c = kc[l];
        if (level < 0 && j < total + c) { level = l; k = j - total; }
        total += c;
    }
    if (j == 0 && lane == 0) counts[f] = min(total, cap);
    if (level < 0 || j >= cap) return;
    const LevelGeom &L = G.lv[level];
    const uint32_t xyr = B.kept[(long long)f * G.keptCap + L.keptOff + k];
    const int x = xyr & 0xfff, y = (xyr >> 12) & 0xfff;
    const float angle = icAngleWarp(G, levelPtr(G, B, f, level) + (long long)y * L.pitch + x, L.pitch, lane);
    const int val = orbByteWarp(B.blur + (long long)f * B.blurSlab + B.blurOff[level] + (long long)y * B.blurPitch[level] + x,
                                B.blurPitch[level], angle, lane);
    const long long o = (long long)f * cap + j;
    desc[o * 32 + lane] = (uint8_t)val;
    if (lane == 0) {
        ygzorb_keypoint kp;
        kp.x = (float)x; kp.y = (float)y;
        if (level != 0) { kp.x = __fmul_rn(kp.x, L.scale); kp.y = __fmul_rn(kp.y, L.scale); }
        kp.size = L.kpSize; kp.angle = angle; kp.response = (float)(xyr >> 24);
        kp.octave = level; kp.class_id = -1;
        kps[o] = kp;
    }
}

int launchDescribe(const FrameGeom &G, const DevBuffers &B, int nFrames, int cap, int32_t *counts,
                   ygzorb_keypoint *kps, uint8_t *desc, cudaStream_t s) {
    const int blocks = (min(cap, G.keptCap) + kDescWarps - 1) / kDescWarps;
    if (B.pdl) launchChain(true, k_describe<true>, dim3(max(blocks, 1), nFrames), dim3(kDescWarps * 32), 0, s, G, B, cap, counts, kps, desc);
    else k_describe<false><<<dim3(max(blocks, 1), nFrames), kDescWarps * 32, 0, s>>>(G, B, cap, counts, kps, desc);
    return 1;
}


// Orientation + descriptor for a given list of patch centres (level, x, y): the descriptor pass of
// operator()(Frame*, ..., DSO_KEYPOINT) (:1079-1113) and the angle refresh of the existing keypoints (:1383-1385).
// A centre closer than 16 px to the border of its level (the reference would read outside the image) gets a zero descriptor
// and keeps its angle.
template <bool kGivenAngle>
__global__ void __launch_bounds__(kDescWarps * 32) k_describe_list(const __grid_constant__ FrameGeom G, const DevBuffers B, int cap,
                                                                   const int4 *centre, const int32_t *counts, ygzorb_keypoint *kps,
                                                                   uint8_t *desc) {
    const int f = blockIdx.y, lane = threadIdx.x & 31;
    const int j = blockIdx.x * kDescWarps + (threadIdx.x >> 5);
    if (j >= min(counts[f], cap)) return;
    const long long o = (long long)f * cap + j;
    const int4 c = centre[o];
    const int level = c.x, x = c.y, y = c.z;
    const LevelGeom &L = G.lv[level];
    if (x < 16 || y < 16 || x >= L.w - 16 || y >= L.h - 16) { desc[o * 32 + lane] = 0; return; }
    // kGivenAngle: the descriptors of pre-existing keys in front of new OctTree keys (:1088-1099) use the angle the key carries
    const float angle = kGivenAngle ? kps[o].angle
                                    : icAngleWarp(G, levelPtr(G, B, f, level) + (long long)y * L.pitch + x, L.pitch, lane);
    const int val = orbByteWarp(B.blur + (long long)f * B.blurSlab + B.blurOff[level] + (long long)y * B.blurPitch[level] + x,
                                B.blurPitch[level], angle, lane);
    desc[o * 32 + lane] = (uint8_t)val;
    if (!kGivenAngle && lane == 0) kps[o].angle = angle;
}

int launchDescribeList(const FrameGeom &G, const DevBuffers &B, int nFrames, int cap, const int4 *centre, const int32_t *counts,
                       ygzorb_keypoint *kps, uint8_t *desc, cudaStream_t s, bool givenAngle) {
    if (nFrames <= 0 || cap <= 0) return 0;
    const dim3 grid((cap + kDescWarps - 1) / kDescWarps, nFrames);
    if (givenAngle) k_describe_list<true><<<grid, kDescWarps * 32, 0, s>>>(G, B, cap, centre, counts, kps, desc);
    else k_describe_list<false><<<grid, kDescWarps * 32, 0, s>>>(G, B, cap, centre, counts, kps, desc);
    return 1;
}

// K0 — cv::remap(src, dst, map1 (CV_16SC2), map2 (CV_16UC1), INTER_LINEAR), BORDER_CONSTANT 0, 8UC1: the undistortion of
// Frame::ComputeImagePyramid (src/Frame.cc:775-805) in front of the pyramid.  Fixed point as OpenCV's remapBilinear: 5
// fractional bits per axis in map2, weights (32-fx)(32-fy)*32 ... summing to 1 << 15, result (sum + (1 << 14)) >> 15; a tap
// outside the source reads the border value 0.  Thread = 4 consecutive destination pixels.
__global__ void __launch_bounds__(256) k_remap(const uint8_t *src, long long srcFrameStride, int srcPitch, int w, int h,
                                               const short2 *map1, const uint16_t *map2, uint8_t *dst) {
    const int x0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y, f = blockIdx.z;
    if (x0 >= w) return;
    const uint8_t *S = src + (long long)f * srcFrameStride;
    uint8_t *D = dst + ((long long)f * h + y) * w;
    const long long mo = (long long)y * w + x0;
    unsigned packed = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (x0 + k >= w) break;
        const short2 m = map1[mo + k];
        const int idx = map2[mo + k] & 1023, fx = idx & 31, fy = idx >> 5;
        const int sx = m.x, sy = m.y;
        int p00 = 0, p01 = 0, p10 = 0, p11 = 0;
        if ((unsigned)sx < (unsigned)(w - 1) && (unsigned)sy < (unsigned)(h - 1)) {
            const uint8_t *r = S + (long long)sy * srcPitch + sx;
            p00 = r[0]; p01 = r[1]; p10 = r[srcPitch]; p11 = r[srcPitch + 1];
        } else {
            const bool xa = (unsigned)sx < (unsigned)w, xb = (unsigned)(sx + 1) < (unsigned)w;
            const bool ya = (unsigned)sy < (unsigned)h, yb = (unsigned)(sy + 1) < (unsigned)h;
            if (ya && xa) p00 = S[(long long)sy * srcPitch + sx];
            if (ya && xb) p01 = S[(long long)sy * srcPitch + sx + 1];
            if (yb && xa) p10 = S[(long long)(sy + 1) * srcPitch + sx];
            if (yb && xb) p11 = S[(long long)(sy + 1) * srcPitch + sx + 1];
        }
        const int acc = ((32 - fx) * p00 + fx * p01) * (32 - fy) + ((32 - fx) * p10 + fx * p11) * fy;   // weights / 32
        packed |= (unsigned)((acc * 32 + (1 << 14)) >> 15) << (8 * k);
    }
    if (x0 + 3 < w && (w & 3) == 0) *reinterpret_cast<unsigned *>(D + x0) = packed;
    else for (int k = 0; k < 4 && x0 + k < w; k++) D[x0 + k] = (uint8_t)(packed >> (8 * k));
}

int launchRemap(const uint8_t *src, long long srcFrameStride, int srcPitch, int w, int h, int nFrames, const short2 *map1,
                const uint16_t *map2, uint8_t *dst, cudaStream_t s) {
    if (nFrames <= 0) return 0;
    k_remap<<<dim3((w + 1023) / 1024, h, nFrames), 256, 0, s>>>(src, srcFrameStride, srcPitch, w, h, map1, map2, dst);
    return 1;
}

static bool gConfigured[64];

int configureKernels() {
    int dev = 0;
    YGZ_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && gConfigured[dev]) return 0;
    YGZ_CUDA(cudaMemcpyToSymbol(gPattern, kOrbPattern, sizeof(kOrbPattern)));
    YGZ_CUDA(cudaFuncSetAttribute(k_quadtree<false, kQtThreads>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    YGZ_CUDA(cudaFuncSetAttribute(k_quadtree<true, kQtThreads>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    YGZ_CUDA(cudaFuncSetAttribute(k_quadtree<false, kQtThreadsBatch>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    if (dev >= 0 && dev < 64) gConfigured[dev] = true;
    return 0;
}

}  // namespace ygz
