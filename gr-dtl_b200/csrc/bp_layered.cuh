// bp_layered.cuh -- layered normalised min-sum decoder, fp32, one codeword per warp (DTL_LDPC_LAYERED: throughput mode).
//
// The decoder BASELINE.json's north_star names next to the reference one: same code tables, same hard-decision exit (zero
// syndrome), different schedule and arithmetic -- it decodes to the same codeword whenever both converge, iteration counts
// differ (DESIGN.md section 4.3 states what is compared).
//
// Schedule.  The check rows are coloured on the host into STEPS of at most 32 rows that share no variable (api.cc:
// schedule_rows: a balanced colouring, 5 steps for the 148 rows of the n = 300 code); one step = one layer, one LANE per row,
// so the updates of a step are independent and the posteriors it reads already contain every earlier step's messages of the
// same iteration (that is what makes the schedule layered).
//
// Layout, all in the warp's shared memory:
//   post[ns + 1]   posteriors, indexed by SLOT: the host assigns every variable a slot (api.cc: lay_assign_slots) and orders
//                  each row's edges so that the j-th access of a step's 32 lanes falls on 32 different banks wherever the graph
//                  allows (per step a bank holds at most RMAX of the step's variables; the edge order is then a bipartite
//                  edge colouring rows x banks).  post[ns] = +1e30 is the pad edge of rows lighter than RMAX.
//   R[steps][32][3] the check-to-variable messages of a row in compressed form: the two smallest normalised magnitudes and a
//                  word with the position of the smallest and the sign bits -- min-sum messages take only two magnitudes per
//                  row.  Stride 3 words per lane: conflict-free.  Not initialised: the first iteration does not read it.
// Per row, all in registers: r_j from the compressed message, q_j = post[v_j] - r_j, the two smallest |q|, the sign parity,
// r_j' = norm * sign * min over the others, post[v_j] = q_j + r_j'.  After every iteration the syndrome of the hard decisions
// is taken straight from the posteriors' signs with the same descriptors (warp vote) and the warp leaves as soon as the word
// is a codeword.
#pragma once
#include "bp_ref.cuh"

namespace dtl {

struct LayS {                       // shared-memory view of one code's tables
    int n, k, rank, ns, max_row, nsteps;
    int o_rows, o_permute;
    const unsigned char* smb;
    __device__ __forceinline__ const uint4* rows() const { return (const uint4*)(smb + o_rows); }   // [nsteps][32]: 7 slots (pad: ns) + row id (0xffff: idle lane)
    __device__ __forceinline__ const uint16_t* permute() const { return (const uint16_t*)(smb + o_permute); }   // H column (cldpc permute applied) of wire position q -> slot
};
struct LayBuf { float* post; uint32_t* r; };

__device__ __forceinline__ LayBuf lay_carve(unsigned char* base, int ns, int steps)
{
    LayBuf w;
    w.r = (uint32_t*)base;
    w.post = (float*)(base + (((size_t)steps * 32 * 3 * 4 + 15) & ~(size_t)15));
    return w;
}
__device__ __forceinline__ LayS stage_lay(const CodeDev& g, unsigned char* sm, int off)
{
    uint4* d = (uint4*)(sm + off);
    uint4* rows = d; d += g.lay_steps * 32;
    uint16_t* perm = (uint16_t*)d;
    for (int i = threadIdx.x; i < g.lay_steps * 32; i += blockDim.x) rows[i] = ((const uint4*)g.lay_rows)[i];
    for (int i = threadIdx.x; i < g.n; i += blockDim.x) perm[i] = g.lay_perm[i];
    LayS c;
    c.n = g.n; c.k = g.k; c.rank = g.rank; c.ns = g.lay_ns; c.max_row = g.max_row; c.nsteps = g.lay_steps;
    c.o_rows = off; c.o_permute = off + 16 * g.lay_steps * 32; c.smb = sm;
    return c;
}

// One iteration = all steps in order.  RMAX = the code's maximum row weight; pad edges of lighter rows point at post[ns],
// which holds +1e30: they never are a minimum and carry a + sign.  FIRST: every message is still zero, R is not read.
template <int RMAX, bool FIRST>
__device__ __forceinline__ void lay_iteration(const LayS& c, const LayBuf& w, int lane, float norm)
{
    const uint4* rows = c.rows() + lane;
    uint32_t* rp = w.r + lane * 3;
    for (int s = 0; s < c.nsteps; s++, rows += 32, rp += 96) {
        int idx[8];
        unpack8(*rows, idx);
        if (idx[7] != 0xffff) {
            float m1o = 0.f, m2o = 0.f;
            unsigned meta = 0;
            if (!FIRST) { m1o = __uint_as_float(rp[0]); m2o = __uint_as_float(rp[1]); meta = rp[2]; }
            const int argo = (int)(meta >> 8);
            float q[RMAX];
            float m1 = 3.0e38f, m2 = 3.0e38f;
            unsigned neg = 0, sgn = 0;
            int arg = 0;
#pragma unroll
            for (int j = 0; j < RMAX; j++) {
                float pv = w.post[idx[j]];
                if (!FIRST) {
                    const float mag = j == argo ? m2o : m1o;
                    pv -= ((meta >> j) & 1u) ? -mag : mag;
                }
                q[j] = pv;
                const float a = fabsf(pv);
                const unsigned sb = __float_as_uint(pv) >> 31;
                neg ^= sb; sgn |= sb << j;
                if (a < m1) { m2 = m1; m1 = a; arg = j; } else m2 = fminf(m2, a);
            }
            m1 *= norm; m2 *= norm;
            const unsigned nsg = neg ? ~sgn : sgn;                  // sign of the new message on edge j: parity of the OTHER edges' signs
#pragma unroll
            for (int j = 0; j < RMAX; j++) {
                if (idx[j] == c.ns) continue;                       // pad edge
                const float mag = j == arg ? m2 : m1;
                w.post[idx[j]] = q[j] + (((nsg >> j) & 1u) ? -mag : mag);
            }
            rp[0] = __float_as_uint(m1); rp[1] = __float_as_uint(m2); rp[2] = (nsg & 0xffu) | ((unsigned)arg << 8);
        }
        __syncwarp();
    }
}

// zero syndrome of the hard decisions post < 0 ?  (same descriptors, one lane per row)
template <int RMAX>
__device__ __forceinline__ bool lay_syndrome(const LayS& c, const LayBuf& w, int lane)
{
    bool bad = false;
    const uint4* rows = c.rows() + lane;
    for (int s = 0; s < c.nsteps; s++, rows += 32) {
        int idx[8];
        unpack8(*rows, idx);
        unsigned par = 0;
#pragma unroll
        for (int j = 0; j < RMAX; j++) par ^= __float_as_uint(w.post[idx[j]]) >> 31;      // pad slot: +1e30, sign 0
        bad |= par != 0;                                                                   // idle lanes: RMAX pad slots
    }
    return __any_sync(0xffffffffu, bad);
}

// post[] holds the received values by slot (rx > 0 => bit 0), post[ns] the pad; returns the iteration count, the decisions are
// the signs of post[].
template <int RMAX>
__device__ __forceinline__ int lay_loop(const LayS& c, const LayBuf& w, int lane, int max_it)
{
    int nit = 0;
    if (max_it > 0) {
        nit = 1;
        lay_iteration<RMAX, true>(c, w, lane, 0.8125f);
        if (!lay_syndrome<RMAX>(c, w, lane)) return nit;
    }
    while (nit < max_it) {
        nit++;
        lay_iteration<RMAX, false>(c, w, lane, 0.8125f);
        if (!lay_syndrome<RMAX>(c, w, lane)) break;
    }
    return nit;
}

__device__ __forceinline__ int bp_lay_run(const LayS& c, const LayBuf& w, int lane, int max_it)
{
    switch (c.max_row) {
    case 3: return lay_loop<3>(c, w, lane, max_it);
    case 4: return lay_loop<4>(c, w, lane, max_it);
    case 5: return lay_loop<5>(c, w, lane, max_it);
    case 6: return lay_loop<6>(c, w, lane, max_it);
    default: return lay_loop<7>(c, w, lane, max_it);
    }
}
__device__ __forceinline__ bool bp_lay_syndrome(const LayS& c, const LayBuf& w, int lane)
{
    switch (c.max_row) {
    case 3: return lay_syndrome<3>(c, w, lane);
    case 4: return lay_syndrome<4>(c, w, lane);
    case 5: return lay_syndrome<5>(c, w, lane);
    case 6: return lay_syndrome<6>(c, w, lane);
    default: return lay_syndrome<7>(c, w, lane);
    }
}

} // namespace dtl
