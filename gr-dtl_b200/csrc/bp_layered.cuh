// bp_layered.cuh -- layered normalised min-sum decoder, fp32, one codeword per warp (DTL_LDPC_LAYERED: throughput mode).
//
// The decoder BASELINE.json's north_star names next to the reference one: same code tables, same hard-decision exit (zero
// syndrome), different schedule and arithmetic -- it decodes to the same codeword whenever both converge, iteration counts
// differ (DESIGN.md section 4 states what is compared).  Schedule: the check rows are coloured on the host into groups that
// share no variable (api.cc: schedule_rows; 13 groups for the n = 300 code); one step = one group, one LANE per row, so the
// updates of a step are independent and the posteriors it reads already contain every earlier group's messages of the
// same iteration (that is what makes the schedule layered).  Per row, all in registers: q_j = post[v_j] - r_j, the two
// smallest |q|, the sign parity, r_j' = norm * sign * min over the others, post[v_j] = q_j + r_j'.  Messages r (m x RS floats) and posteriors (n floats) live in the warp's shared memory: 4.8 KB for
// n = 300, so the occupancy limit is registers, not shared memory.  After every iteration the hard decisions go through the
// same syndrome check as the reference decoder (warp vote) and the warp leaves as soon as the word is a codeword.
#pragma once
#include "bp_ref.cuh"

namespace dtl {

struct LayS {                       // shared-memory view of one code's tables (rv2 / permute as in FastS)
    int n, m, k, rank, rs, max_row, nsteps;
    int o_rv2, o_rows, o_permute;
    const unsigned char* smb;
    __device__ __forceinline__ const uint4* rv2() const { return (const uint4*)(smb + o_rv2); }
    __device__ __forceinline__ const uint4* rows() const { return (const uint4*)(smb + o_rows); }   // [nsteps][32] descriptors
    __device__ __forceinline__ const uint16_t* permute() const { return (const uint16_t*)(smb + o_permute); }
};
struct LayBuf { float* post; float* r; uint8_t* est; };

__device__ __forceinline__ LayBuf lay_carve(unsigned char* base, int n, int m, int rs)
{
    LayBuf w;
    w.r = (float*)base;
    w.post = (float*)(base + (((size_t)m * rs * 4 + 15) & ~(size_t)15));
    w.est = (uint8_t*)(base + (((size_t)m * rs * 4 + 15) & ~(size_t)15) + ((((size_t)n + 2) * 4 + 15) & ~(size_t)15));
    return w;
}
__device__ __forceinline__ LayS stage_lay(const CodeDev& g, unsigned char* sm, int off)
{
    uint4* d = (uint4*)(sm + off);
    uint4* rv2 = d; d += g.m;
    uint4* rows = d; d += g.lay_steps * 32;
    uint16_t* perm = (uint16_t*)d;
    for (int i = threadIdx.x; i < g.m; i += blockDim.x) rv2[i] = g.rv2[i];
    for (int i = threadIdx.x; i < g.lay_steps * 32; i += blockDim.x) rows[i] = ((const uint4*)g.lay_rows)[i];
    for (int i = threadIdx.x; i < g.n; i += blockDim.x) perm[i] = g.permute[i];
    LayS c;
    c.n = g.n; c.m = g.m; c.k = g.k; c.rank = g.rank; c.rs = g.rs; c.max_row = g.max_row; c.nsteps = g.lay_steps;
    c.o_rv2 = off; c.o_rows = off + 16 * g.m; c.o_permute = c.o_rows + 16 * g.lay_steps * 32; c.smb = sm;
    return c;
}

// One iteration = all groups in order.  RMAX = the code's maximum row weight; pad edges of lighter rows point at post[n],
// which holds +inf magnitude (1e30): they never are a minimum and carry a + sign.
template <int RMAX, int RS>
__device__ __forceinline__ void lay_iteration(const LayS& c, const LayBuf& w, int lane, float norm)
{
    for (int s = 0; s < c.nsteps; s++) {
        int idx[8];
        unpack8(c.rows()[s * 32 + lane], idx);
        const int r = idx[7];
        if (r != 0xffff) {
            float* rr = w.r + r * RS;
            float q[RMAX], a[RMAX];
            float m1 = 3.0e38f, m2 = 3.0e38f;
            unsigned neg = 0;
            int arg = 0;
#pragma unroll
            for (int j = 0; j < RMAX; j++) {
                q[j] = w.post[idx[j]] - rr[j];
                a[j] = fabsf(q[j]);
                neg ^= __float_as_uint(q[j]) >> 31;
                if (a[j] < m1) { m2 = m1; m1 = a[j]; arg = j; } else m2 = fminf(m2, a[j]);
            }
            m1 *= norm; m2 *= norm;
#pragma unroll
            for (int j = 0; j < RMAX; j++) {
                if (idx[j] == c.n) continue;                            // pad edge
                const float mag = j == arg ? m2 : m1;
                const float nm = ((neg ^ (__float_as_uint(q[j]) >> 31)) & 1u) ? -mag : mag;
                rr[j] = nm;
                w.post[idx[j]] = q[j] + nm;
            }
        }
        __syncwarp();
    }
}

template <int RMAX>
__device__ __forceinline__ bool lay_syndrome(const LayS& c, const LayBuf& w, int lane)
{
    for (int v = lane; v < c.n; v += 32) w.est[v] = w.post[v] < 0.f ? 1 : 0;
    __syncwarp();
    bool bad = false;
    for (int r = lane; r < c.m; r += 32) {
        int idx[8];
        unpack8(c.rv2()[r], idx);
        unsigned par = 0;
#pragma unroll
        for (int j = 0; j < RMAX; j++) par ^= w.est[idx[j]];
        bad |= (par & 1u) != 0;
    }
    return __any_sync(0xffffffffu, bad);
}

// post[] holds the received values in H column order (rx > 0 => bit 0); returns the iteration count, est[] = decisions.
template <int RMAX, int RS>
__device__ __forceinline__ int lay_loop(const LayS& c, const LayBuf& w, int lane, int max_it)
{
    for (int e = lane; e < c.m * RS; e += 32) w.r[e] = 0.f;
    if (lane == 0) { w.post[c.n] = 1.0e30f; w.est[c.n] = 0; }
    __syncwarp();
    int nit = 0;
    while (nit < max_it) {
        nit++;
        lay_iteration<RMAX, RS>(c, w, lane, 0.8125f);
        if (!lay_syndrome<RMAX>(c, w, lane)) break;
    }
    return nit;
}

__device__ __forceinline__ int bp_lay_run(const LayS& c, const LayBuf& w, int lane, int max_it)
{
    switch (c.max_row) {
    case 3: return lay_loop<3, 4>(c, w, lane, max_it);
    case 4: return lay_loop<4, 4>(c, w, lane, max_it);
    case 5: return lay_loop<5, 6>(c, w, lane, max_it);
    case 6: return lay_loop<6, 6>(c, w, lane, max_it);
    case 7: return lay_loop<7, 8>(c, w, lane, max_it);
    default: return lay_loop<8, 8>(c, w, lane, max_it);
    }
}

} // namespace dtl
