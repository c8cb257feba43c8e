// TEST-SIDE DIAGNOSTIC (uses the CPU oracle, so it lives under tests/): replays the blend traversal of one view on the CPU and
// counts, for alternative lane mappings of the blend kernels, how many serial (warp, Gaussian) visits they would need.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <stdio.h>
// out[]: counters
// 0: instances considered (pos < tile_last)        1: tile-level survivors (any px alpha-pass, ignoring order)
// 2: (inst, 8x4 block) iterations bwd [any px alpha-pass && pos<blocklast]   3: of those with a truly contributing px
// 4: contributing pairs (bwd)    5: half-warp scheme per-32-chunk max iterations (4x4 blocks; warp = 8x4 = L,R)
// 6: half-warp scheme, whole-list max   7: quarter scheme (4x2 blocks, 4 per warp) per-chunk max  8: quarter whole-list max
// 9: sum over (inst, 4x4 block) hits    10: sum over (inst,4x2 block) hits
// 11: fwd (inst, 8x4 block) iterations [any alpha-pass px && block not all done]  12: fwd half per-chunk max 13: fwd contributing pairs
// 14: fwd instances considered (pos < tile's last needed pos)
void blend_stats(int W, int H, const uint32_t* ranges, const uint32_t* plist, const float* means2D,
                 const float* conop, const uint32_t* n_contrib, double* out, int tile_stride, int tile_phase) {
    int gx = (W + 15) / 16, gy = (H + 15) / 16;
    for (int t = tile_phase; t < gx * gy; t += tile_stride) {
        int tx = t % gx, ty = t / gx;
        uint32_t r0 = ranges[2 * t], r1 = ranges[2 * t + 1];
        int n = (int)(r1 - r0);
        if (n <= 0) continue;
        uint32_t lastc[256]; uint32_t tile_last = 0; int inside[256];
        for (int p = 0; p < 256; ++p) {
            int x = tx * 16 + (p & 15), y = ty * 16 + (p >> 4);
            inside[p] = x < W && y < H;
            lastc[p] = inside[p] ? n_contrib[(size_t)y * W + x] : 0;
            if (lastc[p] > tile_last) tile_last = lastc[p];
        }
        int nb = n < (int)tile_last ? n : (int)tile_last;
        // 8x4 blocks: block id = (py/4)*2 + (px/8); 4x4: (py/4)*4 + px/4 ; 4x2: (py/2)*4+px/4
        uint32_t bl84[8] = {0}, bl44[16] = {0}, bl42[32] = {0};
        for (int p = 0; p < 256; ++p) {
            int px = p & 15, py = p >> 4;
            int b84 = (py / 4) * 2 + px / 8, b44 = (py / 4) * 4 + px / 4, b42 = (py / 2) * 4 + px / 4;
            if (lastc[p] > bl84[b84]) bl84[b84] = lastc[p];
            if (lastc[p] > bl44[b44]) bl44[b44] = lastc[p];
            if (lastc[p] > bl42[b42]) bl42[b42] = lastc[p];
        }
        // per-list totals for whole-list max
        long tot44[16] = {0}, tot42[32] = {0};
        long chunk44[16] = {0}, chunk42[32] = {0};
        int surv_in_chunk = 0;
        // fwd state
        float T[256]; int done[256]; int ndone = 0;
        for (int p = 0; p < 256; ++p) { T[p] = 1.0f; done[p] = !inside[p]; ndone += done[p]; }
        long fchunk44[16] = {0}; int fsurv_in_chunk = 0;
        for (int pos = 0; pos < n; ++pos) {
            uint32_t id = plist[r0 + pos];
            float gxm = means2D[2 * id], gym = means2D[2 * id + 1];
            float a = conop[4 * id], b = conop[4 * id + 1], c = conop[4 * id + 2], op = conop[4 * id + 3];
            uint32_t pass[8] = {0};  // 256-bit mask of alpha-passing pixels
            int anypass = 0;
            for (int p = 0; p < 256; ++p) {
                float dx = gxm - (float)(tx * 16 + (p & 15)), dy = gym - (float)(ty * 16 + (p >> 4));
                float power = -0.5f * (a * dx * dx + c * dy * dy) - b * dx * dy;
                if (power > 0.0f) continue;
                float alpha = fminf(0.99f, op * expf(power));
                if (alpha < 1.0f / 255.0f) continue;
                pass[p >> 5] |= 1u << (p & 31); anypass = 1;
            }
            // ---------- forward
            if (ndone < 256) {
                out[14] += 1;
                if (anypass) {
                    int h84[8] = {0}, h44[16] = {0};
                    for (int p = 0; p < 256; ++p) {
                        if (!((pass[p >> 5] >> (p & 31)) & 1)) continue;
                        if (done[p]) continue;
                        int px = p & 15, py = p >> 4;
                        h84[(py / 4) * 2 + px / 8] = 1; h44[(py / 4) * 4 + px / 4] = 1;
                        float dx = gxm - (float)(tx * 16 + px), dy = gym - (float)(ty * 16 + py);
                        float power = -0.5f * (a * dx * dx + c * dy * dy) - b * dx * dy;
                        float alpha = fminf(0.99f, op * expf(power));
                        float tt = T[p] * (1 - alpha);
                        if (tt < 0.0001f) { done[p] = 1; ndone++; continue; }
                        T[p] = tt; out[13] += 1;
                    }
                    int any = 0;
                    for (int k = 0; k < 8; ++k) { out[11] += h84[k]; any |= h84[k]; }
                    for (int k = 0; k < 16; ++k) fchunk44[k] += h44[k];
                    if (any) fsurv_in_chunk++;
                    if (fsurv_in_chunk == 32) {
                        for (int w = 0; w < 8; ++w) { int l = (w / 2) * 4 + (w & 1) * 2; long m = fchunk44[l] > fchunk44[l + 1] ? fchunk44[l] : fchunk44[l + 1]; out[12] += m; }
                        memset(fchunk44, 0, sizeof fchunk44); fsurv_in_chunk = 0;
                    }
                }
            }
            // ---------- backward
            if (pos >= nb) continue;
            out[0] += 1;
            if (!anypass) continue;
            out[1] += 1;
            int h84[8] = {0}, c84[8] = {0}, h44[16] = {0}, h42[32] = {0};
            for (int p = 0; p < 256; ++p) {
                if (!((pass[p >> 5] >> (p & 31)) & 1)) continue;
                int px = p & 15, py = p >> 4;
                int b84 = (py / 4) * 2 + px / 8, b44 = (py / 4) * 4 + px / 4, b42 = (py / 2) * 4 + px / 4;
                if ((uint32_t)pos < bl84[b84]) h84[b84] = 1;
                if ((uint32_t)pos < bl44[b44]) h44[b44] = 1;
                if ((uint32_t)pos < bl42[b42]) h42[b42] = 1;
                if ((uint32_t)pos < lastc[p]) { c84[b84] = 1; out[4] += 1; }
            }
            for (int k = 0; k < 8; ++k) { out[2] += h84[k]; out[3] += c84[k]; }
            for (int k = 0; k < 16; ++k) { out[9] += h44[k]; tot44[k] += h44[k]; chunk44[k] += h44[k]; }
            for (int k = 0; k < 32; ++k) { out[10] += h42[k]; tot42[k] += h42[k]; chunk42[k] += h42[k]; }
            surv_in_chunk++;
            if (surv_in_chunk == 32 || pos == nb - 1) {
                for (int w = 0; w < 8; ++w) {  // warp w owns 8x4 block w: row w/2, col w&1 -> 4x4 blocks (w/2)*4 + (w&1)*2 + {0,1}
                    int l = (w / 2) * 4 + (w & 1) * 2;
                    long m = chunk44[l] > chunk44[l + 1] ? chunk44[l] : chunk44[l + 1];
                    out[5] += m;
                    // 4x2 blocks of the 8x4 block: rows (w/2)*2 + {0,1}, cols (w&1)*2+{0,1}
                    long mm = 0;
                    for (int rr = 0; rr < 2; ++rr) for (int cc = 0; cc < 2; ++cc) {
                        long v = chunk42[((w / 2) * 2 + rr) * 4 + (w & 1) * 2 + cc]; if (v > mm) mm = v; }
                    out[7] += mm;
                }
                memset(chunk44, 0, sizeof chunk44); memset(chunk42, 0, sizeof chunk42); surv_in_chunk = 0;
            }
        }
        if (fsurv_in_chunk) for (int w = 0; w < 8; ++w) { int l = (w / 2) * 4 + (w & 1) * 2; long m = fchunk44[l] > fchunk44[l + 1] ? fchunk44[l] : fchunk44[l + 1]; out[12] += m; }
        for (int w = 0; w < 8; ++w) {
            int l = (w / 2) * 4 + (w & 1) * 2;
            out[6] += tot44[l] > tot44[l + 1] ? tot44[l] : tot44[l + 1];
            long mm = 0;
            for (int rr = 0; rr < 2; ++rr) for (int cc = 0; cc < 2; ++cc) {
                long v = tot42[((w / 2) * 2 + rr) * 4 + (w & 1) * 2 + cc]; if (v > mm) mm = v; }
            out[8] += mm;
        }
    }
}
