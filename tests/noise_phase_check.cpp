// CPU check of the stream arithmetic of the noise phase of the chained launch (fluid_sim_b200/csrc/noise.cuh, chain.cu:
// chain_noise): for every (block, warp, lane) the schedule of polynomial jumps must land on output 2 (row H + c0) of the
// step, the circular generation must produce the outputs of the row piece in order, and the fixed stride jump must
// land on the same columns 32 * kNzWarps rows further down.  Checked against sequential generation from the base window.
#include "noise.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace fsb;

template <int CW>
static int check(int W, int H, unsigned seed, unsigned long long pos)
{
    std::vector<uint32_t> polys((size_t)kNzPolys * 32);
    nz_fill_polys(H, CW, polys.data());
    uint32_t base[kRandDeg];
    rand_seed_window(seed, base);
    if (pos) { RandJump J = jump_matrix(pos); apply_jump(J, base); }
    // the step's outputs, sequentially
    std::vector<uint32_t> seq((size_t)2 * W * H);
    {
        uint32_t w[kRandDeg];
        for (int m = 0; m < kRandDeg; ++m) w[m] = base[m];
        for (size_t n = 0; n < seq.size(); ++n) {
            const uint32_t x = w[0] + w[28];
            for (int j = 0; j < kRandDeg - 1; ++j) w[j] = w[j + 1];
            w[kRandDeg - 1] = x;
            seq[n] = x;
        }
    }
    const int nblk = (H + CW - 1) / CW;
    long long checked = 0;
    for (int blk = 0; blk < nblk; ++blk)
        for (int wid = 0; wid < kNzWarps; ++wid)
            for (int lane = 0; lane < 32; ++lane) {
                // the loop of chain_noise, one thread
                const int c0 = blk * CW;
                uint32_t w[kRandDeg];
                for (int m = 0; m < kRandDeg; ++m) w[m] = base[m];
                int R0 = 32 * wid;
                for (int it = 0;; ++it) {
                    const uint32_t *p;
                    bool take = true;
                    if (it < 10) {
                        if (!((blk >> it) & 1)) continue;
                        p = polys.data() + (kNzBlockPoly + it) * 32;
                    } else if (it < 14) {
                        if (!((wid >> (it - 10)) & 1)) continue;
                        p = polys.data() + (kNzWarpPoly + it - 10) * 32;
                    } else if (it < 19) {
                        take = (lane >> (it - 14)) & 1;
                        p = polys.data() + (kNzLanePoly + it - 14) * 32;
                    } else {
                        if (R0 >= W) break;
                        const int row = R0 + lane;
                        for (int round = 0; round < CW / 32; ++round) {
                            const int rot = (64 * round) % kRandDeg;
                            for (int n = 0; n < 64; ++n) {
                                const uint32_t x = w[(n + rot) % kRandDeg] + w[(n + rot + 28) % kRandDeg];
                                w[(n + rot) % kRandDeg] = x;
                                const int j = c0 + 32 * round + n / 2;
                                if (row < W && j < H) {
                                    const size_t at = 2 * ((size_t)row * H + j) + (n & 1);
                                    if (seq[at] != x) {
                                        printf("W %d H %d cw %d: block %d warp %d lane %d row %d col %d output %d differs\n", W, H, CW,
                                               blk, wid, lane, row, j, n & 1);
                                        return 1;
                                    }
                                    ++checked;
                                }
                            }
                        }
                        R0 += 32 * kNzWarps;
                        if (R0 >= W) break;
                        constexpr int rot_end = (2 * CW) % kRandDeg;
                        uint32_t t[kRandDeg];
                        for (int m = 0; m < kRandDeg; ++m) t[m] = w[(m + rot_end) % kRandDeg];
                        for (int m = 0; m < kRandDeg; ++m) w[m] = t[m];
                        p = polys.data() + kNzStridePoly * 32;
                    }
                    nz_jump(w, p, take);
                }
            }
    if (checked != (long long)2 * W * H) {
        printf("W %d H %d cw %d: %lld outputs checked, %lld expected\n", W, H, CW, checked, (long long)2 * W * H);
        return 1;
    }
    return 0;
}

int main()
{
    if (check<64>(700, 200, 1, 0)) return 1;          // rows beyond one round of the CTA's warps, ragged last block
    if (check<32>(333, 100, 1, 80000)) return 1;
    if (check<64>(40, 1024 + 64, 12345, 7)) return 1;  // block indices with several bits set
    if (check<32>(321, 96, 1, 0)) return 1;
    printf("ok\n");
    return 0;
}
