// fps_multi.cu — multi-pick farthest point sampling for sm_100a (the default FPS kernel).
//
// Replaces sampling/tf_sampling_g.cu:105-170 of the reference; same results bit for bit.
//
// Why: FPS is m-1 strictly sequential arg-max steps.  On a B200 one step of a 16384-point
// cloud is ~130 cycles of arithmetic (the cloud lives in registers, spread over a cluster of
// 8 SMs) and ~1000 cycles of reduction + DSMEM exchange latency (profiles/r1_ncu_full_v5).
// This kernel pays that latency once per ROUND and takes up to K = 8 picks out of it.
//
// Exactness of taking several picks from one arg-max exchange.  Let td[] be the running
// minimum distances after the picks so far and key(j) = (td[j], tie(j)) the reference's total
// order (max distance, then lowest index mod 512, then lowest index; tf_sampling_g.cu:146-163).
// The reference's next pick is r1 = argmax key.  Let r2 be the arg-max over the remaining
// points.  Picking r1 replaces td[j] by min(td[j], d(r1, j)), which never raises a key.  So if
// d(r1, r2) >= td[r2] (r2 is untouched by the update; fminf(d, td) returns td bit for bit) and
// td[r2] > 0 (already-picked points, whose td is 0, stay below it), r2 is still the arg-max
// after the update: it IS the reference's following pick.  By induction r_i is the i-th next
// pick as long as d(r_j, r_i) >= td[r_i] for every j < i.  The first candidate that fails the
// test ends the round; picks it would have made are found by the next round's exchange.
//
// Candidates.  Every warp (a "group") reports the best point it holds, and the value `sec` of
// its second best.  Sorted by key, candidate r_i is the true i-th arg-max unless a point hidden
// behind an already accepted candidate of the same warp beats it, so r_i is also required to
// satisfy td[r_i] > sec(g) for the warps g of r_1..r_{i-1}.  All tests are conservative: a
// failed test only ends the round early, it never changes an index.
//
// Machinery per round (no __syncthreads and no cluster barrier in the loop):
//   1. every thread lowers the td of its PTS register-resident points against the A centres
//      accepted last round (A x PTS x 7 FP32 instructions);
//   2. per thread top-2 + which point; warp arg-max with redux.sync; the owning lane pushes
//      {x, y, z, sec, tie word, td bits} (24 B) into the warp's slot in EVERY CTA of the
//      cluster with st.async (DSMEM), completing on the receiver's mbarrier;
//   3. every warp waits on its CTA's mbarrier, ranks the <= 64 candidates by 64-bit key with
//      one pass of broadcast shared-memory reads, and tests the 28 (i > j) pairs of the top 8
//      on 28 lanes at once; one ballot gives the number A of accepted picks;
//   4. CTA 0 writes the A indices; everyone reads the A centres from the slots next round.
// Measured on 16 x (16384 -> 4096): 6.2 picks per round.
#include "common.cuh"

namespace hf {
namespace {

constexpr int MP_K = 8;         // candidates examined (and at most accepted) per round
constexpr int MP_MSG = 24;      // bytes pushed per warp, per destination CTA, per round
constexpr int MP_SLOT_W = 8;    // 32-bit words reserved per slot (32 B, keeps slots 16 B aligned)

__device__ __forceinline__ uint32_t mp_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void mp_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::
                     : "memory");
}
__device__ __forceinline__ uint32_t mp_mapa(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mp_st_async_v4(uint32_t raddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d,
                                               uint32_t rbar) {
    asm volatile(
        "st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1,%2,%3,%4}, [%5];" ::"r"(
            raddr),
        "r"(a), "r"(b), "r"(c), "r"(d), "r"(rbar)
        : "memory");
}
__device__ __forceinline__ void mp_st_async_v2(uint32_t raddr, uint32_t a, uint32_t b, uint32_t rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1,%2}, [%3];" ::"r"(
                     raddr),
                 "r"(a), "r"(b), "r"(rbar)
                 : "memory");
}
// Low word of the composite key: larger = preferred.  Encodes (k mod 512, k) ascending.
__device__ __forceinline__ uint32_t mp_tie_word(uint32_t k) { return ~(((k & 511u) << 23) | (k >> 9)); }
__device__ __forceinline__ uint32_t mp_tie_decode(uint32_t lo) {
    uint32_t inv = ~lo;
    return ((inv & 0x7FFFFFu) << 9) | (inv >> 23);
}

// Thread -> point mapping.  The cluster's TT = C*T threads cover C*T*PTS points.  Thread slot
// u = g + G*lane (g = the warp's group number rank*W + warp) owns residue(s) r = k mod 512 and
// every R-th block of 512 points:  register p = s*Q + q  is point  k = 512*(u/512 + R*q) + (u%512 + TT*s)
// with S = max(1, 512/TT) residues per thread, R = max(1, TT/512) threads per residue, Q = PTS/S.
//   * a thread's points share k mod 512 (or have ascending residues), so the reference's tie
//     order among them is simply ascending p;
//   * CONSECUTIVE indices live in consecutive groups.  This matters: the deeper levels of a
//     PointNet++ stack sample a cloud that is already in FPS order, where the picks are
//     0, 1, 2, ...; were they in the same warp, its one candidate per round would allow a
//     single pick per round.
// Shared-memory state of one cluster CTA that survives from one FPS level to the next.
template <int G, int W>
struct MpShared {
    uint32_t slots[2][G][MP_SLOT_W];  // {x, y, z, sec | tie word, td bits, -, -}
    uint64_t bars[2];
    uint8_t pos[W][8];  // per warp: rank -> slot of the top 8 candidates
};

// One FPS level (n points -> m picks) run by a whole cluster.  `load(k)` returns point k of the level's cloud
// (global memory for the first level of a launch, the previous level's picks in shared memory afterwards).
//   idxs    (m) global: the picked indices                                         [written by CTA 0]
//   oxyz    (m,3) global or null: the picked points = gather_point(inp, idxs)       [written by CTA 0]
//   nxt     (m,3) shared or null: the same, a full copy in EVERY CTA (the next level's cloud)
// `round` is the cluster-wide exchange round counter: it runs on across the levels of one launch, so the two
// mbarriers keep their phase (each is armed for round + 2 right after it completed for round).
// CH = 0: the coordinates of a thread's PTS points stay in registers.  CH > 0: only the running distances do; the
// coordinates are re-read from the CTA's shared-memory copy every round, CH points at a time (3 conflict-free LDS per
// point and round against 6.5 FP32 instructions per point and accepted centre: hidden behind the arithmetic).  That
// halves the registers per point, so a CTA holds 16384 points (512 threads x 32) instead of 8192 and a 65536-point
// cloud needs 4 SMs instead of 8: 16 scans run in one wave where the register shape needs two (a B200 co-schedules 15
// clusters of 8, tools/gpc_probe.cu).
//
// SORTED: the cloud arrives in the order of its spatial index (spatial.cu: Hilbert-sorted {x, y, z, original index}
// records, `load(s)` = record s, `load.first()` = original point 0) and a WARP holds 32 * PTS consecutive records, i.e.
// a compact region of space with a bounding box.  A centre whose exact lower-bound distance to that box (same operation
// sequence as sqdist, so LB <= fl(d) for every point inside) is not below the warp's current maximum td cannot lower any
// td of the warp (fminf(d, td) == td): the warp skips it.  After the first few hundred picks four centres out of five
// are skipped, and the distance update is the largest part of this kernel's instructions.  The picks are unchanged:
// every candidate carries its ORIGINAL index in the key (shared-memory array behind the coordinates), and ties inside a
// thread or a warp are resolved on that key instead of on the register order.
// SORTED = 2: the compact unit is the THREAD, not the warp: thread slot u holds records [u * PTS, (u+1) * PTS), a bucket
// a metre or two across with its own box and its own maximum td, and consecutive buckets go to consecutive warps, so a
// warp's 32 buckets are scattered over the whole cloud.  That keeps what makes several picks per round possible — a
// warp's candidate is the best of a random-looking sample of the cloud (compact warps give 4.7 instead of 5.3 picks per
// round) — and prunes harder: a lane lists the accepted centres that can reach ITS bucket (a few per cent of them), and
// the warp runs as many update passes as its busiest lane has centres, typically one or two instead of all A.
template <int C, int W, int PTS, int CH = 0, int SORTED = 0, class Load>
__device__ __forceinline__ void fps_multi_level(int n, int m, Load load, int *__restrict__ idxs,
                                                float *__restrict__ oxyz, float *nxt, float *sx,
                                                MpShared<C * W, W> &sh, int &round) {
    constexpr int T = W * 32, TT = C * T, S = TT < 512 ? 512 / TT : 1, R = TT >= 512 ? TT / 512 : 1, Q = PTS / S;
    constexpr int G = C * W, NC = (G + 31) / 32;
    static_assert(PTS % S == 0 && Q >= 1, "PTS must be a multiple of the residues per thread");
    static_assert(TT % 512 == 0 || 512 % TT == 0, "cluster threads must divide or be a multiple of 512");
    static_assert(G <= 64, "at most 64 warps per scene");
    static_assert(PTS <= 32, "the equal-to-max mask is 32 bits");
    constexpr int CAP = T * PTS;  // points held by one CTA
    static_assert(!SORTED || CH == 0, "the sorted shape keeps its coordinates in registers");
    float *sy = sx + CAP;
    float *sz = sy + CAP;
    uint32_t *sorig = reinterpret_cast<uint32_t *>(sz + CAP);  // SORTED: original index of every point of this CTA
    auto &slots = sh.slots;
    auto &bars = sh.bars;
    auto &pos = sh.pos;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t rank = C > 1 ? mp_ctarank() : 0u;
    const uint32_t grp = rank * W + warp;  // this warp's group = its slot number
    // point held in register p of the thread with slot u
    auto kof = [](uint32_t u, int p) -> uint32_t {
        return 512u * (u / 512u + (uint32_t)R * (uint32_t)(p % Q)) + (u % 512u + (uint32_t)TT * (uint32_t)(p / Q));
    };
    const uint32_t u = grp + (uint32_t)G * lane;

    constexpr int XR = CH > 0 ? CH : PTS;  // coordinates held in registers at a time
    static_assert(CH == 0 || PTS % CH == 0, "the chunk must divide the points per thread");
    float x[XR], y[XR], z[XR], td[PTS];
    float blx = INFINITY, bly = INFINITY, blz = INFINITY, bhx = -INFINITY, bhy = -INFINITY, bhz = -INFINITY;  // SORTED: warp box
#pragma unroll
    for (int p = 0; p < PTS; ++p) {
        const int li = p * T + t;  // local copy, conflict free
        float3 v = make_float3(0.f, 0.f, 0.f);
        td[p] = -1.f;  // padding: never a candidate
        if constexpr (SORTED) {
            const uint32_t s_ = SORTED == 2 ? u * (uint32_t)PTS + (uint32_t)p  // a thread = PTS consecutive records
                                            : grp * (32u * PTS) + (uint32_t)p * 32u + lane;  // a warp = 32 * PTS consecutive records
            uint32_t orig = 0u;
            if (s_ < (uint32_t)n) {
                const float4 r = load(s_);
                if (__float_as_int(r.w) >= 0) {
                    v = make_float3(r.x, r.y, r.z);
                    orig = (uint32_t)__float_as_int(r.w);
                    td[p] = 1e38f;
                    blx = fminf(blx, v.x); bly = fminf(bly, v.y); blz = fminf(blz, v.z);
                    bhx = fmaxf(bhx, v.x); bhy = fmaxf(bhy, v.y); bhz = fmaxf(bhz, v.z);
                }
            }
            sorig[li] = orig;
        } else {
            const uint32_t k = kof(u, p);
            if (k < (uint32_t)n) {
                v = load(k);
                td[p] = 1e38f;  // tf_sampling_g.cu:118
            }
        }
        if (CH == 0) {
            x[p] = v.x;
            y[p] = v.y;
            z[p] = v.z;
        }
        sx[li] = v.x;
        sy[li] = v.y;
        sz[li] = v.z;
    }
    if constexpr (SORTED == 1) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            blx = fminf(blx, __shfl_xor_sync(0xffffffffu, blx, o)); bly = fminf(bly, __shfl_xor_sync(0xffffffffu, bly, o));
            blz = fminf(blz, __shfl_xor_sync(0xffffffffu, blz, o)); bhx = fmaxf(bhx, __shfl_xor_sync(0xffffffffu, bhx, o));
            bhy = fmaxf(bhy, __shfl_xor_sync(0xffffffffu, bhy, o)); bhz = fmaxf(bhz, __shfl_xor_sync(0xffffffffu, bhz, o));
        }
    }
    // original index of the point in register p of lane l_ of this warp (tie words are built from it)
    auto orig_of = [&](int l_, int p) -> uint32_t {
        if constexpr (SORTED) return sorig[p * T + warp * 32 + l_];
        else return kof(grp + (uint32_t)G * (uint32_t)l_, p);
    };
    uint32_t wmax_prev = 0x7F800000u;  // SORTED = 1: bits of the warp's maximum td after the previous round
    float tmax_prev = INFINITY;        // SORTED = 2: this thread's maximum td after the previous round
    if (lane < 8) pos[warp][lane] = 0;
    float3 first;  // every thread reads it BEFORE the barrier: `nxt` may alias the cloud `load` reads
    if constexpr (SORTED) first = load.first();
    else first = load(0);
    __syncthreads();
    if (t == 0) {
        // round 0 of the level reads its single centre (point 0, tf_sampling_g.cu:114-116) like any other round
        uint32_t *c0 = slots[(round & 1) ^ 1][0];
        c0[0] = __float_as_uint(first.x);
        c0[1] = __float_as_uint(first.y);
        c0[2] = __float_as_uint(first.z);
        if (nxt) {
            nxt[0] = first.x;
            nxt[1] = first.y;
            nxt[2] = first.z;
        }
        if (rank == 0) {
            idxs[0] = 0;
            if (oxyz) {
                oxyz[0] = first.x;
                oxyz[1] = first.y;
                oxyz[2] = first.z;
            }
        }
    }
    __syncthreads();

    // lane c < C pushes to CTA c: remote address of this warp's slot (buffer 0) and of barrier 0 there
    const uint32_t rslot = mp_mapa(smem_u32(&slots[0][grp][0]), (uint32_t)(lane % C));
    const uint32_t rbar = mp_mapa(smem_u32(&bars[0]), (uint32_t)(lane % C));
    // lane q < 28 tests candidate pair (pi, pj), pi > pj, q = pi*(pi-1)/2 + pj
    int pi = 1, pj = 0;
    for (int q = 0; q < lane && q < 27; ++q) {
        if (++pj == pi) {
            ++pi;
            pj = 0;
        }
    }
    const int sh_i = 8 * pi, sh_j = 8 * pj;

    uint64_t posw = 0;  // the 8 bytes of pos[warp]: slot of the candidate of rank 0..7
    int A = 1, done = 1;
    for (; done < m; ++round) {
        const int buf = round & 1;
        const uint32_t(*cs)[MP_SLOT_W] = slots[buf ^ 1];  // centres accepted last round
        // ---- 1. lower td against the A new centres, two at a time (min is exact in any order) ----
        if constexpr (SORTED == 2) {
            unsigned todo = 0u;  // accepted centres that can reach this thread's bucket
            for (int i = 0; i < A; ++i) {
                const float4 a = *reinterpret_cast<const float4 *>(cs[(uint32_t)(posw >> (8 * i)) & 0xffu]);
                const float gx = fmaxf(0.f, fmaxf(blx - a.x, a.x - bhx));
                const float gy = fmaxf(0.f, fmaxf(bly - a.y, a.y - bhy));
                const float gz = fmaxf(0.f, fmaxf(blz - a.z, a.z - bhz));
                todo |= sqdist(gx, gy, gz) < tmax_prev ? (1u << i) : 0u;  // exact lower bound, as for the warp box
            }
            while (__any_sync(0xffffffffu, todo != 0u)) {
                if (todo) {
                    const int i = __ffs(todo) - 1;
                    todo &= todo - 1u;
                    const float4 a = *reinterpret_cast<const float4 *>(cs[(uint32_t)(posw >> (8 * i)) & 0xffu]);
#pragma unroll
                    for (int p = 0; p < PTS; ++p) td[p] = fminf(sqdist(x[p] - a.x, y[p] - a.y, z[p] - a.z), td[p]);
                }
            }
        } else if constexpr (SORTED == 1) {
            const float wmaxf = __uint_as_float(wmax_prev);
            for (int i = 0; i < A; ++i) {
                const float4 a = *reinterpret_cast<const float4 *>(cs[(uint32_t)(posw >> (8 * i)) & 0xffu]);
                // exact lower bound of sqdist(point - a) over the warp's box (same operation sequence as sqdist)
                const float gx = fmaxf(0.f, fmaxf(blx - a.x, a.x - bhx));
                const float gy = fmaxf(0.f, fmaxf(bly - a.y, a.y - bhy));
                const float gz = fmaxf(0.f, fmaxf(blz - a.z, a.z - bhz));
                if (sqdist(gx, gy, gz) < wmaxf) {  // warp uniform
#pragma unroll
                    for (int p = 0; p < PTS; ++p) td[p] = fminf(sqdist(x[p] - a.x, y[p] - a.y, z[p] - a.z), td[p]);
                }
            }
        } else if constexpr (CH > 0) {
#pragma unroll
            for (int c0 = 0; c0 < PTS; c0 += CH) {
#pragma unroll
                for (int q = 0; q < CH; ++q) {
                    const int li = (c0 + q) * T + t;
                    x[q] = sx[li];
                    y[q] = sy[li];
                    z[q] = sz[li];
                }
                int i = 0;
                for (; i + 2 <= A; i += 2) {
                    const float4 a = *reinterpret_cast<const float4 *>(cs[(uint32_t)(posw >> (8 * i)) & 0xffu]);
                    const float4 b2 = *reinterpret_cast<const float4 *>(cs[(uint32_t)(posw >> (8 * (i + 1))) & 0xffu]);
#pragma unroll
                    for (int q = 0; q < CH; ++q)
                        td[c0 + q] = fminf(fminf(sqdist(x[q] - a.x, y[q] - a.y, z[q] - a.z),
                                                 sqdist(x[q] - b2.x, y[q] - b2.y, z[q] - b2.z)),
                                           td[c0 + q]);
                }
                if (i < A) {
                    const float4 a = *reinterpret_cast<const float4 *>(cs[(uint32_t)(posw >> (8 * i)) & 0xffu]);
#pragma unroll
                    for (int q = 0; q < CH; ++q) td[c0 + q] = fminf(sqdist(x[q] - a.x, y[q] - a.y, z[q] - a.z), td[c0 + q]);
                }
            }
        } else {
            int i = 0;
            float4 c0 = *reinterpret_cast<const float4 *>(cs[(uint32_t)posw & 0xffu]);
            float4 c1 = *reinterpret_cast<const float4 *>(cs[(uint32_t)(posw >> 8) & 0xffu]);
            for (; i + 2 <= A; i += 2) {
                const float4 a = c0, b2 = c1;
                // prefetch the next pair (ranks i+2, i+3 <= 7; stale bytes beyond A are valid slot numbers)
                c0 = *reinterpret_cast<const float4 *>(cs[(uint32_t)(posw >> (8 * ((i + 2) & 7))) & 0xffu]);
                c1 = *reinterpret_cast<const float4 *>(cs[(uint32_t)(posw >> (8 * ((i + 3) & 7))) & 0xffu]);
#pragma unroll
                for (int p = 0; p < PTS; ++p)
                    td[p] = fminf(fminf(sqdist(x[p] - a.x, y[p] - a.y, z[p] - a.z),
                                        sqdist(x[p] - b2.x, y[p] - b2.y, z[p] - b2.z)),
                                  td[p]);
            }
            if (i < A) {
#pragma unroll
                for (int p = 0; p < PTS; ++p) td[p] = fminf(sqdist(x[p] - c0.x, y[p] - c0.y, z[p] - c0.z), td[p]);
            }
        }
        // ---- 2. this thread's best point and the value of its second best ----
        float m1 = td[0];
#pragma unroll
        for (int p = 1; p < PTS; ++p) m1 = fmaxf(m1, td[p]);
        tmax_prev = m1;               // (padding only: -1, no centre passes the test)
        const bool have = m1 >= 0.f;  // false: only padding here
        const uint32_t hi = have ? __float_as_uint(m1) : 0u;
        const uint32_t wm = __reduce_max_sync(0xffffffffu, hi);  // warp maximum (in flight during the scan below)
        wmax_prev = wm;
        uint32_t eqm = 0u;
        float m2a = -1.f, m2b = -1.f;
#pragma unroll
        for (int p = 0; p < PTS; ++p) {
            const bool e = td[p] == m1;
            eqm += e ? (1u << p) : 0u;
            if (p & 1) m2b = fmaxf(m2b, e ? -1.f : td[p]);
            else m2a = fmaxf(m2a, e ? -1.f : td[p]);
        }
        float m2 = fmaxf(m2a, m2b);
        if (eqm & (eqm - 1u)) m2 = m1;  // two points share the maximum
        int bp = __ffs(eqm) - 1;  // first in the reference's tie order (registers ascend in it unless SORTED)
        if constexpr (SORTED) {
            if (eqm & (eqm - 1u)) {  // rare (duplicate points): the reference's order is on the ORIGINAL index
                uint32_t best = 0u;
                for (uint32_t rest = eqm; rest; rest &= rest - 1u) {
                    const int p = __ffs(rest) - 1;
                    const uint32_t w_ = mp_tie_word(sorig[p * T + t]);
                    if (w_ >= best) {
                        best = w_;
                        bp = p;
                    }
                }
            }
        }
        const uint32_t h2 = m2 >= 0.f ? __float_as_uint(m2) : 0u;
        const bool top = have && hi == wm;
        const unsigned eq = __ballot_sync(0xffffffffu, top);
        uint32_t ws = __reduce_max_sync(0xffffffffu, top ? h2 : hi);  // second best of the warp
        int L = __ffs(eq) - 1;
        if (eq & (eq - 1u)) {  // several lanes hold the warp maximum (duplicate points): rare, warp-uniform
            ws = wm;
            const uint32_t mylo = mp_tie_word(orig_of(lane, bp));
            const uint32_t wl = __reduce_max_sync(0xffffffffu, top ? mylo : 0u);
            L = __ffs(__ballot_sync(0xffffffffu, top && mylo == wl)) - 1;
        }
        // ---- push the warp's candidate: lane c sends it to CTA c ----
        {
            const int wl_ = L < 0 ? 0 : L;
            const int wbp = __shfl_sync(0xffffffffu, bp, wl_);
            if (lane < C) {
                uint32_t mx = 0u, my = 0u, mz = 0u, mlo = 0u, mhi = 0u;
                if (L >= 0) {
                    const int wli = wbp * T + warp * 32 + wl_;
                    mx = __float_as_uint(sx[wli]);
                    my = __float_as_uint(sy[wli]);
                    mz = __float_as_uint(sz[wli]);
                    mlo = mp_tie_word(orig_of(wl_, wbp));
                    mhi = wm;
                }
                const uint32_t so = (uint32_t)buf * (G * MP_SLOT_W * 4), bo = (uint32_t)buf * 8;
                mp_st_async_v4(rslot + so, mx, my, mz, ws, rbar + bo);
                mp_st_async_v2(rslot + so + 16, mlo, mhi, rbar + bo);
            }
        }
        // ---- 3. wait for all G candidates ----
        mbar_wait(&bars[buf], (uint32_t)(round >> 1) & 1u);
        if (t == 0) mbar_arrive_expect_tx(&bars[buf], G * MP_MSG);  // for round + 2
        const uint32_t(*sl)[MP_SLOT_W] = slots[buf];
        uint64_t key[NC];
        uint32_t rk[NC][4];
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const int s = lane + 32 * c;
            key[c] = s < G ? *reinterpret_cast<const uint64_t *>(&sl[s][4]) : 0ull;
            rk[c][0] = rk[c][1] = rk[c][2] = rk[c][3] = 0u;
        }
#pragma unroll
        for (int s = 0; s < G; ++s) {
            const uint64_t ks = *reinterpret_cast<const uint64_t *>(&sl[s][4]);
#pragma unroll
            for (int c = 0; c < NC; ++c) rk[c][s & 3] += ks > key[c] ? 1u : 0u;  // 4 independent chains
        }
        __syncwarp();
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const uint32_t r = (rk[c][0] + rk[c][1]) + (rk[c][2] + rk[c][3]);
            if (lane + 32 * c < G && r < (uint32_t)MP_K) pos[warp][r] = (uint8_t)(lane + 32 * c);
        }
        __syncwarp();
        posw = *reinterpret_cast<const uint64_t *>(&pos[warp][0]);
        // ---- pair tests: does candidate pi survive the acceptance of candidate pj? (branch free) ----
        const uint32_t *ca = sl[(uint32_t)(posw >> sh_i) & 0xffu];
        const uint32_t *cb = sl[(uint32_t)(posw >> sh_j) & 0xffu];
        const float4 va = *reinterpret_cast<const float4 *>(ca);
        const float4 vb = *reinterpret_cast<const float4 *>(cb);
        const uint32_t hia = ca[5];
        const float dab = sqdist(va.x - vb.x, va.y - vb.y, va.z - vb.z);
        const bool viol = (pi >= G) | (hia == 0u) | (__float_as_uint(vb.w) >= hia) | (dab < __uint_as_float(hia));
        const unsigned vm = __ballot_sync(0xffffffffu, lane < 28 && viol);
        {
            // first violating pair f (lanes are ordered by pi) -> pi(f) = number of accepted candidates
            const int f = __ffs(vm) - 1;
            const int row = 1 + (f >= 1) + (f >= 3) + (f >= 6) + (f >= 10) + (f >= 15) + (f >= 21);
            A = vm ? row : MP_K;
            A = min(A, m - done);
        }
        // ---- 4. emit: index and coordinates (the slot carries them) of the A accepted picks ----
        if (warp == 0 && lane < A) {
            const uint32_t *w = sl[(uint32_t)(posw >> (8 * lane)) & 0xffu];
            if (nxt) {
                float *o = nxt + 3 * (done + lane);
                o[0] = __uint_as_float(w[0]);
                o[1] = __uint_as_float(w[1]);
                o[2] = __uint_as_float(w[2]);
            }
            if (rank == 0) {
                idxs[done + lane] = (int)mp_tie_decode(w[4]);
                if (oxyz) {
                    float *o = oxyz + 3 * (size_t)(done + lane);
                    o[0] = __uint_as_float(w[0]);
                    o[1] = __uint_as_float(w[1]);
                    o[2] = __uint_as_float(w[2]);
                }
            }
        }
        done += A;
    }
}

// Set-up shared by the kernels below: slots cleared, both mbarriers armed for their first round.
template <int C, int W>
__device__ __forceinline__ void mp_setup(MpShared<C * W, W> &sh) {
    constexpr int G = C * W, T = W * 32;
    const int t = threadIdx.x;
    for (int i = t; i < 2 * G * MP_SLOT_W; i += T) (&sh.slots[0][0][0])[i] = 0u;
    if (t == 0) {
        mbar_init(&sh.bars[0], 1);
        mbar_init(&sh.bars[1], 1);
        mbar_fence_init();
        mbar_arrive_expect_tx(&sh.bars[0], G * MP_MSG);
        mbar_arrive_expect_tx(&sh.bars[1], G * MP_MSG);
    }
    __syncthreads();
    if (C > 1) mp_cluster_sync();  // nobody pushes into a CTA whose barriers are not armed yet
}

// One level: farthest_point_sample (+ gather_point when new_xyz != null).
template <int C, int W, int PTS, int CH = 0>
__global__ void __launch_bounds__(W * 32, 1)
    fps_multi_kernel(int n, int m, const float *__restrict__ inp, int *__restrict__ out, float *__restrict__ new_xyz) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ __align__(16) MpShared<C * W, W> sh;
    const int scene = blockIdx.x / C;
    const float *data = inp + (size_t)scene * n * 3;
    mp_setup<C, W>(sh);
    int round = 0;
    fps_multi_level<C, W, PTS, CH>(
        n, m, [data](uint32_t k) { return make_float3(data[(size_t)k * 3], data[(size_t)k * 3 + 1], data[(size_t)k * 3 + 2]); },
        out + (size_t)scene * m, new_xyz ? new_xyz + (size_t)scene * m * 3 : nullptr, nullptr,
        reinterpret_cast<float *>(smem_raw), sh, round);
    if (C > 1) mp_cluster_sync();  // no CTA may exit while peers can still push into it
}

// One level over a cloud that comes with its spatial index (SORTED mode of fps_multi_level).
struct SortedLoad {
    const float4 *rec;  // the scene's Hilbert-sorted records {x, y, z, original index | -1 = padding}
    const float *data;  // the scene's points in original order (the first centre is original point 0)
    __device__ __forceinline__ float4 operator()(uint32_t s) const { return __ldg(&rec[s]); }
    __device__ __forceinline__ float3 first() const { return make_float3(data[0], data[1], data[2]); }
};

template <int C, int W, int PTS, int MODE>
__global__ void __launch_bounds__(W * 32, 1)
    fps_multi_sorted_kernel(int n, int npad, int m, const float *__restrict__ inp, const float4 *__restrict__ sorted,
                            int *__restrict__ out, float *__restrict__ new_xyz) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ __align__(16) MpShared<C * W, W> sh;
    const int scene = blockIdx.x / C;
    mp_setup<C, W>(sh);
    int round = 0;
    SortedLoad load{sorted + (size_t)scene * npad, inp + (size_t)scene * n * 3};
    fps_multi_level<C, W, PTS, 0, MODE>(npad, m, load, out + (size_t)scene * m,
                                        new_xyz ? new_xyz + (size_t)scene * m * 3 : nullptr, nullptr,
                                        reinterpret_cast<float *>(smem_raw), sh, round);
    if (C > 1) mp_cluster_sync();  // no CTA may exit while peers can still push into it
}

// The lower levels of a sampling pyramid in ONE launch: level j samples the picks of level j-1, which every CTA of
// the cluster already holds in shared memory (each accepted centre reaches every CTA through the exchange), so
// neither the clouds nor the indices go through HBM between levels and the launches + gathers between them
// disappear (hf/core/feature_extractors/pointnet_util.py:42-44 runs farthest_point_sample + gather_point per
// SA level; levels 2..4 of the RPN chain are 4096 -> 1024 -> 256 -> 64).  Level j keeps PTS >> 2j points per thread.
struct PyramidArgs {
    int levels;     // 1..3
    int n;          // points of the input cloud
    int m[3];       // picks per level, m[j] <= points of level j
    const float *inp;
    int *idx[3];    // (b, m[j]) each
    float *xyz[3];  // (b, m[j], 3) each
};

template <int C, int W, int P0>
__global__ void __launch_bounds__(W * 32, 1) fps_pyramid_kernel(PyramidArgs a) {
    constexpr int P1 = P0 >= 4 ? P0 / 4 : 1, P2 = P1 >= 4 ? P1 / 4 : 1;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ __align__(16) MpShared<C * W, W> sh;
    float *sx = reinterpret_cast<float *>(smem_raw);       // 3 * T * P0 floats: this CTA's part of the level's cloud
    float *cloud = sx + 3 * W * 32 * P0;                   // 3 * m[0] floats: the picks of the level = the next cloud
    const int scene = blockIdx.x / C;
    const float *data = a.inp + (size_t)scene * a.n * 3;
    mp_setup<C, W>(sh);
    int round = 0;
    fps_multi_level<C, W, P0>(
        a.n, a.m[0], [data](uint32_t k) { return make_float3(data[(size_t)k * 3], data[(size_t)k * 3 + 1], data[(size_t)k * 3 + 2]); },
        a.idx[0] + (size_t)scene * a.m[0], a.xyz[0] + (size_t)scene * a.m[0] * 3, a.levels > 1 ? cloud : nullptr, sx, sh, round);
    auto from_smem = [cloud](uint32_t k) { return make_float3(cloud[3 * k], cloud[3 * k + 1], cloud[3 * k + 2]); };
    if (a.levels > 1) {
        if (C > 1) mp_cluster_sync();  // every CTA is out of the level: its slots and staging arrays can be reused
        else __syncthreads();
        fps_multi_level<C, W, P1>(a.m[0], a.m[1], from_smem, a.idx[1] + (size_t)scene * a.m[1],
                                  a.xyz[1] + (size_t)scene * a.m[1] * 3, a.levels > 2 ? cloud : nullptr, sx, sh, round);
    }
    if (a.levels > 2) {
        if (C > 1) mp_cluster_sync();
        else __syncthreads();
        fps_multi_level<C, W, P2>(a.m[1], a.m[2], from_smem, a.idx[2] + (size_t)scene * a.m[2],
                                  a.xyz[2] + (size_t)scene * a.m[2] * 3, nullptr, sx, sh, round);
    }
    if (C > 1) mp_cluster_sync();  // no CTA may exit while peers can still push into it
}

template <class Kern, class... Args>
int launch_cluster(Kern kern, const char *name, int b, int C, int threads, size_t smem, cudaStream_t stream, Args... args) {
    if (smem > 40 * 1024) HF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)b * C);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = C;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    HF_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, args...));
    HF_LAUNCH_OK(name);
    return HF_OK;
}

template <int C, int W, int PTS, int CH = 0>
int launch_fps_multi(int b, int n, int m, const float *inp, int *out, float *new_xyz, cudaStream_t stream) {
    return launch_cluster(fps_multi_kernel<C, W, PTS, CH>, "fps_multi_kernel", b, C, W * 32, (size_t)3 * W * 32 * PTS * sizeof(float),
                          stream, n, m, inp, out, new_xyz);
}

template <int C, int W, int PTS>
int launch_fps_sorted(int mode, int b, int n, int npad, int m, const float *inp, const float4 *sorted, int *out,
                      float *new_xyz, cudaStream_t stream) {
    const size_t smem = (size_t)4 * W * 32 * PTS * sizeof(float);
    if (mode == 1)
        return launch_cluster(fps_multi_sorted_kernel<C, W, PTS, 1>, "fps_multi_sorted_kernel", b, C, W * 32, smem, stream, n,
                              npad, m, inp, sorted, out, new_xyz);
    return launch_cluster(fps_multi_sorted_kernel<C, W, PTS, 2>, "fps_multi_sorted_kernel", b, C, W * 32, smem, stream, n,
                          npad, m, inp, sorted, out, new_xyz);
}

template <int C, int W, int P0>
int launch_fps_pyramid(int b, const PyramidArgs &a, cudaStream_t stream) {
    const size_t smem = ((size_t)3 * W * 32 * P0 + (size_t)3 * a.m[0]) * sizeof(float);
    return launch_cluster(fps_pyramid_kernel<C, W, P0>, "fps_pyramid_kernel", b, C, W * 32, smem, stream, a);
}

template <int C>
int dispatch_multi(int per, int warps, int b, int n, int m, const float *inp, int *out, float *oxyz, cudaStream_t s) {
    // warps per CTA: 4 by default (one per scheduler: the shortest latency chain); 8 / 16 put the same
    // number of groups on half / a quarter of the SMs (hf_fps_set_variant(3, warps, 0)).
    if constexpr (C <= 4) {
      if (warps == 8 && per <= 4096) {
        if (per <= 2048) return launch_fps_multi<C, 8, 8>(b, n, m, inp, out, oxyz, s);
        return launch_fps_multi<C, 8, 16>(b, n, m, inp, out, oxyz, s);
      }
    }
    if constexpr (C <= 2) {
      if (warps == 16 && per <= 8192) {
        if (per <= 4096) return launch_fps_multi<C, 16, 8>(b, n, m, inp, out, oxyz, s);
        return launch_fps_multi<C, 16, 16>(b, n, m, inp, out, oxyz, s);
      }
    }
    if (per <= 512) return launch_fps_multi<C, 4, 4>(b, n, m, inp, out, oxyz, s);
    if (per <= 1024) return launch_fps_multi<C, 4, 8>(b, n, m, inp, out, oxyz, s);
    if (per <= 2048) return launch_fps_multi<C, 4, 16>(b, n, m, inp, out, oxyz, s);
    if (per <= 4096) return launch_fps_multi<C, 4, 32>(b, n, m, inp, out, oxyz, s);
    if (per <= 8192) return launch_fps_multi<C, 8, 32>(b, n, m, inp, out, oxyz, s);
    return HF_ERR_UNSUPPORTED;
}

}  // namespace

// Launch the multi-pick FPS.  cluster_req = 0 -> measured heuristic (tools/fps_multi_sweep.py):
// 1 CTA up to 512 points, 2 up to 2048, 4 up to 4096, else 8; halved while the batch would not
// fit the GPU in one wave and a CTA still holds <= 4096 points.  Returns HF_ERR_UNSUPPORTED
// for clouds beyond 8 x 8192 points.
int fps_multi_dispatch(int b, int n, int m, const float *inp, int *out, float *oxyz, cudaStream_t s, int cluster_req, int warps_req) {
    int cluster = 1;
    if (cluster_req == 1 || cluster_req == 2 || cluster_req == 4 || cluster_req == 8) {
        cluster = cluster_req;
    } else {
        cluster = n <= 512 ? 1 : n <= 2048 ? 2 : n <= 4096 ? 4 : 8;
        const int sms = num_sms();
        while (cluster > 1 && (long long)b * cluster > sms && (n + cluster / 2 - 1) / (cluster / 2) <= 4096) cluster >>= 1;
        // A B200 co-schedules 15 clusters of 8 CTAs, not 18 (a cluster must sit inside one GPC:
        // profiles/r1_fps_wave_probe.txt): from the 16th scene on, clusters share SMs and every
        // scene runs at two warps per scheduler anyway (0.83 -> 1.18 ms for 16384 -> 4096).  Then
        // 4 CTAs of 8 warps hold the same 32 groups on half the SMs for 5% more time alone, and
        // leave the other half of the GPU to concurrent kernels (chain: 2.76 -> 2.56 ms/step).
        if (cluster == 8 && warps_req == 0 && b > 15 && (n + 3) / 4 <= 4096) {
            cluster = 4;
            warps_req = 8;
        }
    }
    // 6 CTAs x 8 warps x 11 points: 16 scans of 16384 points need 96 SMs and 16 clusters of 6 fit the GPCs of a
    // B200 at one CTA per SM, which 16 clusters of 8 do not (tools/gpc_probe.cu).  Opt-in: hf_fps_set_cluster(6).
    if (cluster_req == 6) {
        if (n > 6 * 256 * 11) return HF_ERR_UNSUPPORTED;
        if (n <= 6 * 256 * 3) return launch_fps_multi<6, 8, 3>(b, n, m, inp, out, oxyz, s);
        return launch_fps_multi<6, 8, 11>(b, n, m, inp, out, oxyz, s);
    }
    // Clouds of 16385..65536 points, more than 15 scans: 4 CTAs x 16 warps x 32 points with the coordinates in shared
    // memory (196 KB per CTA), 4 SMs per scan.  The register shape (8 CTAs x 8 warps x 32 points) is twice as fast per
    // scan (8.8 vs 16.4 ms for 65536 -> 16384: 16 warps per SM rank the 64 candidates redundantly and share the FP32
    // pipe) but a B200 holds only 15 clusters of 8, so the 16th scan costs a second wave (16.5 ms): the same time on
    // 64 SMs instead of all of them, the rest stays free for the kernels that run beside the sampling chain
    // (profiles/r2_fps_65536_shapes.txt).  warps_req = 16 / 8 force the shared-memory / register shape.
    if (n > 4 * 4096 && n <= 4 * 16384 && (cluster_req == 0 || cluster_req == 4) &&
        (warps_req == 16 || (warps_req == 0 && b > 15)))
        return launch_fps_multi<4, 16, 32, 8>(b, n, m, inp, out, oxyz, s);
    while (cluster < 8 && (n + cluster - 1) / cluster > 8192) cluster <<= 1;
    const int per = (n + cluster - 1) / cluster;
    switch (cluster) {
        case 1: return dispatch_multi<1>(per, warps_req, b, n, m, inp, out, oxyz, s);
        case 2: return dispatch_multi<2>(per, warps_req, b, n, m, inp, out, oxyz, s);
        case 4: return dispatch_multi<4>(per, warps_req, b, n, m, inp, out, oxyz, s);
        default: return dispatch_multi<8>(per, warps_req, b, n, m, inp, out, oxyz, s);
    }
}


// FPS over a cloud with its spatial index (hf_cloud_index_build): same picks, warps skip the centres that cannot
// reach them.  HF_ERR_UNSUPPORTED when no shape holds the cloud: the caller falls back to the unsorted kernel.
int fps_sorted_dispatch(int b, int n, int m, const float *inp, const void *index, int *out, float *oxyz, cudaStream_t s,
                        int warps_req, int mode_req) {
    const int mode = mode_req == 1 ? 1 : 2;  // 1 = compact warps, 2 = compact threads (default)
    CloudIndexView ix;  // layout owned by spatial.cu
    if (!cloud_index_view(b, n, index, &ix)) return HF_ERR_UNSUPPORTED;
    const int npad = ix.npad;
    const float4 *sorted = ix.sorted;
    if (npad <= 4096) return HF_ERR_UNSUPPORTED;  // small clouds: the exchange, not the update, is the cost
    if (npad <= 16384) {
        if (warps_req == 4 || (warps_req == 0 && b <= 15)) return launch_fps_sorted<8, 4, 16>(mode, b, n, npad, m, inp, sorted, out, oxyz, s);
        // more than 15 scans: 4 CTAs x 8 warps (1.16 ms for 16 x 16384 -> 4096 against 1.25 without the index).  6 CTAs
        // x 8 warps x 11 records (warps_req = 6; 22 such clusters are co-resident) measured 1.49 ms: not the default.
        if (warps_req == 6) return launch_fps_sorted<6, 8, 11>(mode, b, n, npad, m, inp, sorted, out, oxyz, s);
        return launch_fps_sorted<4, 8, 16>(mode, b, n, npad, m, inp, sorted, out, oxyz, s);
    }
    if (npad <= 65536) return launch_fps_sorted<8, 8, 32>(mode, b, n, npad, m, inp, sorted, out, oxyz, s);
    return HF_ERR_UNSUPPORTED;
}

// Levels 2.. of a sampling pyramid in one launch (fps_pyramid_kernel).  HF_ERR_UNSUPPORTED when no instantiation
// holds the shape: the caller falls back to one launch per level.
int fps_pyramid_dispatch(int b, int levels, int n, const int *m, const float *inp, int *const *idx, float *const *xyz,
                         cudaStream_t s) {
    if (levels < 1 || levels > 3 || n > 16384) return HF_ERR_UNSUPPORTED;
    PyramidArgs a = {};
    a.levels = levels;
    a.n = n;
    a.inp = inp;
    for (int j = 0; j < levels; ++j) {
        a.m[j] = m[j];
        a.idx[j] = idx[j];
        a.xyz[j] = xyz[j];
    }
    int tt, p0;
    if (n <= 4096) {
        tt = 512;
        p0 = n <= 512 ? 1 : n <= 1024 ? 2 : n <= 2048 ? 4 : 8;
    } else {
        tt = 1024;
        p0 = 16;
    }
    const int p1 = p0 >= 4 ? p0 / 4 : 1, p2 = p1 >= 4 ? p1 / 4 : 1;
    if (levels > 1 && m[0] > tt * p1) return HF_ERR_UNSUPPORTED;
    if (levels > 2 && m[1] > tt * p2) return HF_ERR_UNSUPPORTED;
    if (m[0] > 8192) return HF_ERR_UNSUPPORTED;  // the picks of the first level are staged in shared memory
    if (tt == 1024) return b > 15 ? launch_fps_pyramid<4, 8, 16>(b, a, s) : launch_fps_pyramid<8, 4, 16>(b, a, s);
    switch (p0) {
        case 1: return launch_fps_pyramid<4, 4, 1>(b, a, s);
        case 2: return launch_fps_pyramid<4, 4, 2>(b, a, s);
        case 4: return launch_fps_pyramid<4, 4, 4>(b, a, s);
        default: return launch_fps_pyramid<4, 4, 8>(b, a, s);
    }
}

}  // namespace hf
