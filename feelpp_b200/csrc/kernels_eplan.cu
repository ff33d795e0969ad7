// kernels_eplan.cu -- entity-task assembly kernel (see eplan.h for the design) and the plan's device mirror.
//
// Replaces, for constant-coefficient mass and stiffness forms on P1/P2 tetrahedra, the reference's per-element
// update/integrate/assemble + MatSetValues(ADD_VALUES) (feelvf/integrator.hpp:1093-1207,
// feelalg/matrixpetsc.cpp:917-936): every CSR value is accumulated in the registers of one thread over star(sigma)
// and stored once, in a summation order fixed by the plan.
#include <string>
#include "feelgpu_internal.h"
#include "eplan.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <memory>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace {

// out_o = sum_k coef[type][o][k] * (sum over the task's elements of the k-th loaded S entry)   (stiffness)
// out_o = mass[type][o] * (sum over the task's elements of c J)                                  (mass)
// ghost rows land directly in their owners' memory: the row-table base of such a row is EPLAN_REMOTE_FLAG | neighbour << 56 |
// offset (in doubles) inside that neighbour's landing buffer ptr[neighbour]
struct ERemote {
    int n, accumulate;
    double* ptr[EPLAN_MAXREMOTE];
};

struct ECoef {
    double coef[EPLAN_MAXTYPES][EPLAN_MAXOUT][EPLAN_MAXLD];
    double mass[EPLAN_MAXTYPES][EPLAN_MAXOUT];
};
__constant__ ECoef c_ecoef;

__host__ __device__ constexpr int e_nld(int order, int type)
{
    return order == 1 ? 1 : (type == 0 ? 1 : type == 1 ? 2 : type == 2 ? 2 : type == 3 ? 3 : type == 4 ? 5 : 4);
}
__host__ __device__ constexpr int e_nout(int order, int type)
{
    return order == 1 ? 1 : (type == 0 ? 1 : type == 1 ? 2 : type == 2 ? 1 : type == 3 ? 3 : type == 4 ? 3 : 1);
}

__device__ __forceinline__ double lds64(uint32_t addr)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
template <int BYTES>
__device__ __forceinline__ void cp_async(void* dst_shared, const void* src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"((uint32_t)__cvta_generic_to_shared(dst_shared)), "l"(src),
                 "n"(BYTES) : "memory");
}
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }
// the same at shared address 0x10000 + off (the S array of the aligned layout; the constant rides in the instruction)
__device__ __forceinline__ double lds64_s(uint32_t off)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1+65536];" : "=d"(v) : "r"(off));
    return v;
}
__device__ __forceinline__ void sts64(uint32_t addr, double v)
{
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}

// one chunk: lane = task.  wp points at this lane's first pair of unit words; words are [ceil(niter/2)][32] pairs (two
// consecutive iterations side by side; the first pair and the out-buffer positions were fetched while the previous chunk
// ran).  A unit word holds the visit slot times 8 in bits 3-11 and the 4-bit S indices from bit 12.
template <int ORDER, int FORM, int TYPE, bool AL>
__device__ __forceinline__ void run_chunk(const uint2* __restrict__ wp, int npair, uint2 cur,
                                          const uint32_t (&pos)[EPLAN_MAXOUT], uint32_t sS, uint32_t vsb, uint32_t sO)
{
    constexpr int NLD = FORM == 1 ? e_nld(ORDER, TYPE) : 1;
    constexpr int NOUT = e_nout(ORDER, TYPE);
    double sum[NLD];
#pragma unroll
    for (int k = 0; k < NLD; ++k) sum[k] = 0.0;
    auto visit = [&](uint32_t w) {
        if (AL) {
            // S starts at shared address 0x10000 (checked by the caller) and holds 4 KB per index: the address of load k is
            // a bit-field composition, the constant rides in the instruction
            const uint32_t va = w & 0xFF8u;
#pragma unroll
            for (int k = 0; k < NLD; ++k) sum[k] += lds64_s((((k == 0 || FORM == 0) ? w : (w >> (4 * k))) & (FORM == 0 ? 0u : 0xF000u)) | va);
        } else {
            const uint32_t va = sS + (w & 0xFF8u);
            if (FORM == 0) {
                sum[0] += lds64(va);
            } else {
#pragma unroll
                for (int k = 0; k < NLD; ++k) sum[k] += lds64(va + ((w >> (EPLAN_IDX_SHIFT + 4 * k)) & 15u) * vsb);
            }
        }
    };
    for (int t = 0; t < npair; ++t) {
        uint2 nxt = make_uint2(0u, 0u);
        if (t + 1 < npair) nxt = __ldg(wp + (t + 1) * 32);
        visit(cur.x);
        visit(cur.y);      // (an odd chunk ends with a word of the all-zero record)
        cur = nxt;
    }
#pragma unroll
    for (int o = 0; o < NOUT; ++o) {
        double v;
        if (FORM == 0) {
            v = c_ecoef.mass[TYPE][o] * sum[0];
        } else {
            v = c_ecoef.coef[TYPE][o][0] * sum[0];
#pragma unroll
            for (int k = 1; k < NLD; ++k) v = fma(c_ecoef.coef[TYPE][o][k], sum[k], v);
        }
        sts64(sO + (pos[o] << 3), v);
    }
}

template <int ORDER, int FORM, int NCOMP, bool AL>   // FORM 0: mass, 1: stiffness; NCOMP components (block = value * Id);
                                                    // AL: the S array sits on a 64 KB boundary of shared memory
__global__ void __launch_bounds__(EPLAN_WARPS * 32, EPLAN_MINB)
k_eplan(const EPlanCta* __restrict__ ctas, int ncta, int* __restrict__ sched, const int32_t* __restrict__ verts, const uint32_t* __restrict__ visits,
        const uint2* __restrict__ chunks, const uint32_t* __restrict__ blob, const uint4* __restrict__ rows,
        const double* __restrict__ xyz, double* __restrict__ a, int accumulate, int vs, int omax, double cval,
        const ERemote rem)
{
    static_assert(EPLAN_WARPS * 32 >= EPLAN_VS - 1 && EPLAN_WARPS * 32 >= EPLAN_ROWS && EPLAN_WARPS * 32 >= EPLAN_VERTS &&
                      EPLAN_WARPS * 32 >= EPLAN_HDRS, "one thread per visit / row / vertex / cached chunk header");
    constexpr int NSREC = FORM == 1 ? 10 : 1;
    extern __shared__ double sm[];
    double* sO = sm;                          // [omax]: the patch's row block; slot 0 = trash
    uint2* sH = reinterpret_cast<uint2*>(sO + omax);   // [EPLAN_HDRS] chunk headers
    double* sS = reinterpret_cast<double*>(sH + EPLAN_HDRS);   // [NSREC][vs] (row block + headers = 63 KB: see AL)
    double* sX = sS + NSREC * vs;             // [3][EPLAN_VERTS]
    int* sCnt = reinterpret_cast<int*>(sX + 3 * EPLAN_VERTS);   // the chunk counter
    uint4* sN = reinterpret_cast<uint4*>(sCnt + 4);    // [2][2] headers of the current and of the next patch (by parity)
    int32_t* sVid = reinterpret_cast<int32_t*>(sN + 4);   // [EPLAN_VERTS] vertex ids of the next patch
    double* sZ = reinterpret_cast<double*>(sVid + EPLAN_VERTS);   // [EPLAN_VERTS] z coordinates of the next patch (staging)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // The CTA is persistent: after patch blockIdx.x it takes patches from a global counter (sched[0]).  Everything a patch's
    // prologue reads from global memory (vertex coordinates and visit record: one of each per thread; row table and chunk
    // headers: asynchronous copies to shared memory) is requested while the previous patch's phase C drains, so no patch
    // but the first pays for the chain header -> vertex id -> coordinates.
    double px = 0.0, py = 0.0;
    uint4 vq = make_uint4(0u, 0u, 0u, 0u);
    auto fetch = [&](const uint4& h0, const uint4& h1, int32_t id) {
        if (tid < (int)(h1.x & 0xFFFF)) {
            const double2 p0 = __ldg(reinterpret_cast<const double2*>(xyz + 4 * (int64_t)id));
            px = p0.x; py = p0.y;
            cp_async<8>(&sZ[tid], xyz + 4 * (int64_t)id + 2);   // (registers are short while phase C runs)
        }
        if (tid < (int)(h1.x >> 16)) {
            if (ORDER == 2) vq = __ldg(reinterpret_cast<const uint4*>(visits + h0.y) + tid);
            else vq.x = __ldg(&visits[h0.y + tid]);
        }
        // chunk headers go straight to shared memory (phase B of the previous patch is over)
        if (tid < (int)(h1.y & 0xFFFF) && tid < EPLAN_HDRS) cp_async<8>(&sH[tid], &chunks[h0.z + tid]);
    };
    auto prefetch_range = [&](const void* p0, uint32_t nbytes, int l0) {
        const uintptr_t lo = reinterpret_cast<uintptr_t>(p0) & ~(uintptr_t)15;
        const uintptr_t hi = (reinterpret_cast<uintptr_t>(p0) + nbytes + 15) & ~(uintptr_t)15;
        for (uintptr_t q = lo + (uintptr_t)l0 * 4096; q < hi; q += 32 * 4096) {
            const uint32_t bytes = (uint32_t)min((uintptr_t)4096, hi - q);
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(q), "r"(bytes) : "memory");
        }
    };
    {
        // vertex lists have a fixed stride, so the id of "my" vertex is requested together with the patch header
        int32_t vid = 0;
        if (tid < EPLAN_VERTS) vid = __ldg(&verts[(size_t)blockIdx.x * EPLAN_VERTS + tid]);
        const uint4* cq = reinterpret_cast<const uint4*>(ctas + blockIdx.x);
        const uint4 c0 = __ldg(cq), c1 = __ldg(cq + 1);
        if (warp == 0) prefetch_range(reinterpret_cast<const uint4*>(blob) + c1.w, (c1.z >> 16) * 16u, lane);
        fetch(c0, c1, vid);
        if (tid == 0) { sN[0] = c0; sN[1] = c1; }
    }
    if (tid < NSREC) sS[tid * vs + vs - 1] = 0.0;             // the all-zero visit record (padding of the chunks)
    __syncthreads();

    for (int par = 0;; par ^= 1) {
        const uint32_t chunk_off = sN[2 * par].z;
        const uint4 c1 = sN[2 * par + 1];
        const int nverts = c1.x & 0xFFFF, nvisit = c1.x >> 16, nchunk = c1.y & 0xFFFF, nrows = c1.y >> 16;
        const uint32_t nout = c1.z & 0x7FFF, needzero = c1.z & 0x8000;
        // ---- phase A0: the prologue data waiting in registers goes to shared memory ----
        cp_async_wait();                       // the chunk headers and z
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the copy engine has read the previous row block
        if (tid < nverts) {
            sX[tid] = px;
            sX[EPLAN_VERTS + tid] = py;
            sX[2 * EPLAN_VERTS + tid] = sZ[tid];
        }
        if (tid == 0) *sCnt = EPLAN_WARPS;     // chunks 0..W-1 are the warps' first ones, the rest is handed out on demand
        if (needzero) {   // only when the graph holds entries that no element produces (owned rows of a partition's interface)
            // every thread has waited for ITS OWN bulk copies of the previous patch only: the row block may be overwritten once
            // all of them have (found by the multi-GPU parity block: 2-3 rows of 43 M came out zero, different ones every run)
            __syncthreads();
            for (uint32_t i = tid; i < (nout + 1) / 2; i += blockDim.x) reinterpret_cast<double2*>(sO)[i] = make_double2(0.0, 0.0);
        }
        __syncthreads();
        // ---- phase A1: geometry of every element touched by this patch's rows, once (thread = visit) ----
        if (tid < nvisit) {
            const int v = tid;
            const uint32_t w = vq.x;
            double x[4][3];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int lv = (w >> (8 * k)) & 255;
                x[k][0] = sX[lv];
                x[k][1] = sX[EPLAN_VERTS + lv];
                x[k][2] = sX[2 * EPLAN_VERTS + lv];
            }
            // K = G grad(N) (edge vectors / 2), det, J = |det| (feelpoly/geomap.hpp:2320-2355, 1971-1976)
            double K[3][3];
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                K[r][0] = 0.5 * (x[1][r] - x[0][r]);
                K[r][1] = 0.5 * (x[2][r] - x[0][r]);
                K[r][2] = 0.5 * (x[3][r] - x[0][r]);
            }
            const double c00 = K[1][1] * K[2][2] - K[1][2] * K[2][1];
            const double c01 = K[1][2] * K[2][0] - K[1][0] * K[2][2];
            const double c02 = K[1][0] * K[2][1] - K[1][1] * K[2][0];
            const double det = K[0][0] * c00 + K[0][1] * c01 + K[0][2] * c02;
            const double cj = cval * fabs(det);
            if (FORM == 0) {
                sS[v] = cj;
                if (ORDER == 2) {   // (edge, opposite edge) entries: one element each, stored here (0 = not a row of this CTA)
                    const double m = c_ecoef.mass[5][0] * cj;
                    sO[vq.y & 0xFFFF] = m; sO[vq.y >> 16] = m; sO[vq.z & 0xFFFF] = m;
                    sO[vq.z >> 16] = m;    sO[vq.w & 0xFFFF] = m; sO[vq.w >> 16] = m;
                }
            } else {
                const double id = 1.0 / det;
                // B = K^{-T} (geomap.hpp:1985-1991): B[r][b] = d xi_b / d x_r ; grad lambda_{b+1} = B[:,b]/2
                double h[4][3];
                h[1][0] = 0.5 * c00 * id;
                h[2][0] = 0.5 * c01 * id;
                h[3][0] = 0.5 * c02 * id;
                h[1][1] = 0.5 * (K[0][2] * K[2][1] - K[0][1] * K[2][2]) * id;
                h[2][1] = 0.5 * (K[0][0] * K[2][2] - K[0][2] * K[2][0]) * id;
                h[3][1] = 0.5 * (K[0][1] * K[2][0] - K[0][0] * K[2][1]) * id;
                h[1][2] = 0.5 * (K[0][1] * K[1][2] - K[0][2] * K[1][1]) * id;
                h[2][2] = 0.5 * (K[0][2] * K[1][0] - K[0][0] * K[1][2]) * id;
                h[3][2] = 0.5 * (K[0][0] * K[1][1] - K[0][1] * K[1][0]) * id;
#pragma unroll
                for (int r = 0; r < 3; ++r) h[0][r] = -(h[1][r] + h[2][r] + h[3][r]);
                // S_kl = c J grad(l_k).grad(l_l): diagonals 0..3, then (0,1),(0,2),(0,3),(1,2),(1,3),(2,3)
#pragma unroll
                for (int k = 0; k < 4; ++k) sS[k * vs + v] = cj * (h[k][0] * h[k][0] + h[k][1] * h[k][1] + h[k][2] * h[k][2]);
                const double s01 = cj * (h[0][0] * h[1][0] + h[0][1] * h[1][1] + h[0][2] * h[1][2]);
                const double s02 = cj * (h[0][0] * h[2][0] + h[0][1] * h[2][1] + h[0][2] * h[2][2]);
                const double s03 = cj * (h[0][0] * h[3][0] + h[0][1] * h[3][1] + h[0][2] * h[3][2]);
                const double s12 = cj * (h[1][0] * h[2][0] + h[1][1] * h[2][1] + h[1][2] * h[2][2]);
                const double s13 = cj * (h[1][0] * h[3][0] + h[1][1] * h[3][1] + h[1][2] * h[3][2]);
                const double s23 = cj * (h[2][0] * h[3][0] + h[2][1] * h[3][1] + h[2][2] * h[3][2]);
                sS[4 * vs + v] = s01;
                sS[5 * vs + v] = s02;
                sS[6 * vs + v] = s03;
                sS[7 * vs + v] = s12;
                sS[8 * vs + v] = s13;
                sS[9 * vs + v] = s23;
                if (ORDER == 2) {
                    // (edge, opposite edge) entries have sigma = the whole tet, i.e. this element alone: value =
                    // x (S_ac + S_ad + S_bc + S_bd) for edges (a,b) | (c,d); local edges e0=(1,2)|e3=(0,3), e1=(2,0)|e4=(1,3),
                    // e2=(0,1)|e5=(2,3) (feelmesh/simplexordering.hpp:283-291).  Slot per row edge; 0 = not a row of this CTA.
                    const double x5 = c_ecoef.coef[5][0][0];
                    const double t0 = x5 * ((s01 + s13) + (s02 + s23));
                    const double t1 = x5 * ((s12 + s23) + (s01 + s03));
                    const double t2 = x5 * ((s02 + s03) + (s12 + s13));
                    sO[vq.y & 0xFFFF] = t0; sO[vq.y >> 16] = t1; sO[vq.z & 0xFFFF] = t2;
                    sO[vq.z >> 16] = t0;    sO[vq.w & 0xFFFF] = t1; sO[vq.w >> 16] = t2;
                }
            }
        }
        __syncthreads();

        // the vertex cache is dead now: it becomes the patch's row table (read in phase C, after the next barrier)
        uint4* sR = reinterpret_cast<uint4*>(sX);
        if (tid < nrows) cp_async<16>(&sR[tid], &rows[sN[2 * par].w + tid]);

        // ---- next patch: before it joins phase B the last warp takes the next patch id from the global counter, pulls that
        // patch's data into L2 (the plan is streamed once from HBM) and leaves its header and vertex ids in shared memory.
        // Nobody waits for this chain of three loads: chunks are handed out on demand and the result is needed after phase B.
        if (warp == EPLAN_WARPS - 1) {
            int pn = 0;
            if (lane == 0) pn = (int)gridDim.x + atomicAdd(sched, 1);
            pn = __shfl_sync(0xffffffffu, pn, 0);
            if (pn < ncta) {
                const uint4* nq = reinterpret_cast<const uint4*>(ctas + pn);
                uint4 n0 = __ldg(nq);
                const uint4 n1 = __ldg(nq + 1);
                int32_t ids[EPLAN_VERTS / 32];
#pragma unroll
                for (int k = 0; k < EPLAN_VERTS / 32; ++k) ids[k] = __ldg(&verts[(size_t)pn * EPLAN_VERTS + 32 * k + lane]);
                prefetch_range(visits + n0.y, (n1.x >> 16) * (ORDER == 2 ? 16u : 4u), lane);
                prefetch_range(chunks + n0.z, (n1.y & 0xFFFF) * 8u, lane);
                prefetch_range(rows + n0.w, (n1.y >> 16) * 16u, lane);
                prefetch_range(reinterpret_cast<const uint4*>(blob) + n1.w, (n1.z >> 16) * 16u, lane);
#pragma unroll
                for (int k = 0; k < EPLAN_VERTS / 32; ++k) {
                    sVid[32 * k + lane] = ids[k];
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(xyz + 4 * (int64_t)ids[k]));
                }
                n0.x = (uint32_t)pn;           // (the vertex offset is implied by the fixed stride)
                if (lane == 0) { sN[2 * (par ^ 1)] = n0; sN[2 * (par ^ 1) + 1] = n1; }
            } else if (lane == 0)
                sN[2 * (par ^ 1)].x = (uint32_t)pn;
        }
        // ---- phase B: task chunks, lane = task.  Chunks are sorted by cost and handed out on demand (shared counter); the next
        // chunk's header, first words and positions are fetched while the current one runs.
        {
            const uint32_t sSa = (uint32_t)__cvta_generic_to_shared(sS), sOa = (uint32_t)__cvta_generic_to_shared(sO);
            if (AL && sSa != 0x10000u) __trap();   // the host checked the layout
            const uint32_t vsb = (uint32_t)vs * 8u, sCa = (uint32_t)__cvta_generic_to_shared(sCnt);
            auto header = [&](int c) { return c < EPLAN_HDRS ? sH[c] : __ldg(&chunks[chunk_off + c]); };
            int c = warp;
            if (c < nchunk) {
                uint2 ch = header(c);
                const uint2* wp = reinterpret_cast<const uint2*>(blob + (size_t)ch.x * 4) + lane;
                uint32_t meta = ch.y;
                uint2 w = __ldg(wp);
                uint32_t pos[EPLAN_MAXOUT];
                {
                    const uint16_t* op = reinterpret_cast<const uint16_t*>(wp - lane + ((((meta >> 8) & 0xFF) + 1) >> 1) * 32) + lane;
#pragma unroll
                    for (int o = 0; o < EPLAN_MAXOUT; ++o) pos[o] = __ldg(op + o * 32);
                }
                while (true) {
                    const int type = meta & 0xFF, npair = (((meta >> 8) & 0xFF) + 1) >> 1;
                    const uint2* wpc = wp;
                    int cn;   // next chunk: one shared-memory atomic by lane 0 (written out: the compiler's warp-aggregated form costs 20 instructions)
                    asm volatile("{ .reg .pred p; setp.eq.u32 p, %1, 0; mov.u32 %0, 0; @p atom.shared.inc.u32 %0, [%2], 0x7fffffff; }"
                                 : "=r"(cn) : "r"(lane), "r"(sCa) : "memory");
                    cn = __shfl_sync(0xffffffffu, cn, 0);
                    const bool more = cn < nchunk;
                    ch = header(more ? cn : c);                  // past the end: re-read the current chunk (harmless)
                    wp = reinterpret_cast<const uint2*>(blob + (size_t)ch.x * 4) + lane;
                    meta = ch.y;
                    const uint2 nw = __ldg(wp);
                    uint32_t npos[EPLAN_MAXOUT];
                    {
                        const uint16_t* op = reinterpret_cast<const uint16_t*>(wp - lane + ((((meta >> 8) & 0xFF) + 1) >> 1) * 32) + lane;
#pragma unroll
                        for (int o = 0; o < EPLAN_MAXOUT; ++o) npos[o] = __ldg(op + o * 32);
                    }
                    switch (type) {
                    case 0: run_chunk<ORDER, FORM, 0, AL>(wpc, npair, w, pos, sSa, vsb, sOa); break;
                    case 1: run_chunk<ORDER, FORM, 1, AL>(wpc, npair, w, pos, sSa, vsb, sOa); break;
                    case 2: if (ORDER >= 2) run_chunk<ORDER, FORM, 2, AL>(wpc, npair, w, pos, sSa, vsb, sOa); break;
                    case 3: if (ORDER >= 2) run_chunk<ORDER, FORM, 3, AL>(wpc, npair, w, pos, sSa, vsb, sOa); break;
                    default: if (ORDER >= 2) run_chunk<ORDER, FORM, 4, AL>(wpc, npair, w, pos, sSa, vsb, sOa); break;   // type 5: phase A
                    }
                    if (!more) break;
                    c = cn;
                    w = nw;
#pragma unroll
                    for (int o = 0; o < EPLAN_MAXOUT; ++o) pos[o] = npos[o];
                }
            }
        }
        if (NCOMP == 1) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // phase C reads the row block through the copy engine
        cp_async_wait();   // the row table
        __syncthreads();

        // the next patch's prologue loads fly while phase C runs
        const bool has_next = (int)sN[2 * (par ^ 1)].x < ncta;
        if (has_next) fetch(sN[2 * (par ^ 1)], sN[2 * (par ^ 1) + 1], tid < EPLAN_VERTS ? sVid[tid] : 0);

        // ---- phase C: the patch's row block goes to the CSR value array, every value once ----
        // a ghost row goes over NVLink into the buffer its owner exported (ranges are ordered: the first is the lowest)
        auto resolve = [&](int64_t base, int& acc) {
            acc = accumulate;
            if (base & EPLAN_REMOTE_FLAG) {
                acc = rem.accumulate;
                return rem.ptr[(base >> 56) & (EPLAN_MAXREMOTE - 1)] + (base & ((int64_t(1) << 56) - 1));
            }
            return a + base;
        };
        if (NCOMP == 1) {
            // Scalar spaces: thread = row segment.  A segment that starts on the same 16-byte phase in the row block as in
            // the CSR array (the host plan arranges it for all but a few) is one bulk copy shared -> global (TMA; a bulk
            // reduction for '+='), plus at most one value by hand at either end; the copy engine reads the row block while
            // the CTA goes on.  The rest is copied by the warp.
            double* dst = nullptr;
            uint32_t soff = 0, len = 0;
            int acc = 0;
            bool byhand = false;
            if (tid < nrows) {
                const uint4 q = sR[tid];
                dst = resolve((int64_t)(((uint64_t)q.y << 32) | q.x), acc);
                soff = q.z;
                len = q.w;
                const uint32_t dphase = (uint32_t)(reinterpret_cast<uintptr_t>(dst) >> 3) & 1u;
                if (dphase == (soff & 1u)) {
                    const uint32_t head = dphase, nmid = (len - head) & ~1u;     // len >= 1
                    if (head) { if (acc) dst[0] += sO[soff]; else dst[0] = sO[soff]; }
                    if (nmid) {
                        const uint32_t sa = (uint32_t)__cvta_generic_to_shared(sO + soff + head);
                        if (acc)
                            asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f64 [%0], [%1], %2;"
                                         ::"l"(dst + head), "r"(sa), "r"(nmid * 8u) : "memory");
                        else
                            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                                         ::"l"(dst + head), "r"(sa), "r"(nmid * 8u) : "memory");
                    }
                    if ((len - head) & 1u) { if (acc) dst[len - 1] += sO[soff + len - 1]; else dst[len - 1] = sO[soff + len - 1]; }
                } else
                    byhand = true;
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            unsigned todo = __ballot_sync(0xffffffffu, byhand);
            while (todo) {
                const int l = __ffs(todo) - 1;
                todo &= todo - 1;
                double* d = reinterpret_cast<double*>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(dst), l));
                const uint32_t so = __shfl_sync(0xffffffffu, soff, l), n = __shfl_sync(0xffffffffu, len, l);
                const int ac = __shfl_sync(0xffffffffu, acc, l);
                for (uint32_t i = lane; i < n; i += 32) { if (ac) d[i] += sO[so + i]; else d[i] = sO[so + i]; }
            }
        } else {
            for (int r = warp; r < nrows; r += EPLAN_WARPS) {
                const uint4 q = sR[r];
                const double* src = sO + q.z;
                const int len = (int)q.w;
                int acc;
                double* rowp = resolve((int64_t)(((uint64_t)q.y << 32) | q.x), acc);
                // product space, block = v * Id: component row c1 starts at base + c1*len*NCOMP
#pragma unroll 1
                for (int i = lane; i < len; i += 32) {
                    const double v = src[i];
#pragma unroll
                    for (int c1 = 0; c1 < NCOMP; ++c1)
#pragma unroll
                        for (int c2 = 0; c2 < NCOMP; ++c2) {
                            double* dst = rowp + (int64_t)c1 * len * NCOMP + (int64_t)i * NCOMP + c2;
                            if (c1 == c2) { if (acc) *dst += v; else *dst = v; }
                            else if (!acc) *dst = 0.0;
                        }
                }
            }
        }
        if (!has_next) break;
        __syncthreads();   // phase C has read the row table and the row block
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    // the last CTA to leave rearms the counters for the next launch
    if (tid == 0 && atomicAdd(sched + 1, 1) == (int)gridDim.x - 1) {
        sched[0] = 0;
        sched[1] = 0;
    }
}

// shared-window address at which a kernel's dynamic shared memory starts (the same for every kernel without static
// shared memory): decides whether the S array of k_eplan falls on a 64 KB boundary
__global__ void k_smem_base(uint32_t* out)
{
    extern __shared__ double sm[];
    *out = (uint32_t)__cvta_generic_to_shared(sm);
}

int smem_base(feelgpu_context_t ctx, cudaStream_t s)
{
    static int cached[64];
    static bool have[64];
    const int d = ctx->device & 63;
    if (!have[d]) {
        uint32_t* dv = nullptr;
        uint32_t hv = 0xFFFFFFFFu;
        if (cudaMalloc(&dv, sizeof(uint32_t)) == cudaSuccess) {
            k_smem_base<<<1, 1, 64, s>>>(dv);
            if (cudaMemcpyAsync(&hv, dv, sizeof(uint32_t), cudaMemcpyDeviceToHost, s) != cudaSuccess ||
                cudaStreamSynchronize(s) != cudaSuccess)
                hv = 0xFFFFFFFFu;
            cudaFree(dv);
        }
        cached[d] = (int)hv;
        have[d] = true;
    }
    return cached[d];
}

size_t eplan_smem_bytes(const EPlan* p, int form)
{
    const int nsrec = form == 1 ? 10 : 1;
    return sizeof(double) * ((size_t)nsrec * EPLAN_VS + 3 * EPLAN_VERTS + (size_t)p->omax) + EPLAN_HDRS * sizeof(uint2) + 16 + 4 * sizeof(uint4) + EPLAN_VERTS * (sizeof(int32_t) + sizeof(double));
}

template <typename T, typename A>
cudaError_t upload(T** d, const std::vector<T, A>& h, int64_t& bytes)
{
    cudaError_t err = cudaMalloc(d, std::max<size_t>(h.size(), 1) * sizeof(T));
    if (err != cudaSuccess) return err;
    bytes += (int64_t)(h.size() * sizeof(T));
    return cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
}

// least squares x of U x = t (U: 6 x m, column-major in u[k][6]); returns the relative residual
double lsq6(int m, const double u[][6], const double* t, double* x)
{
    double A[EPLAN_MAXLD][EPLAN_MAXLD + 1];
    for (int r = 0; r < m; ++r) {
        for (int c = 0; c < m; ++c) {
            double s = 0;
            for (int k = 0; k < 6; ++k) s += u[r][k] * u[c][k];
            A[r][c] = s;
        }
        double s = 0;
        for (int k = 0; k < 6; ++k) s += u[r][k] * t[k];
        A[r][m] = s;
    }
    for (int c = 0; c < m; ++c) {
        int piv = c;
        for (int r = c + 1; r < m; ++r)
            if (std::fabs(A[r][c]) > std::fabs(A[piv][c])) piv = r;
        if (std::fabs(A[piv][c]) < 1e-300) return 1.0;
        for (int k = 0; k <= m; ++k) std::swap(A[c][k], A[piv][k]);
        for (int r = 0; r < m; ++r) {
            if (r == c) continue;
            const double f = A[r][c] / A[c][c];
            for (int k = c; k <= m; ++k) A[r][k] -= f * A[c][k];
        }
    }
    for (int c = 0; c < m; ++c) x[c] = A[c][m] / A[c][c];
    double res = 0, nt = 0;
    for (int k = 0; k < 6; ++k) {
        double s = -t[k];
        for (int c = 0; c < m; ++c) s += u[c][k] * x[c];
        res += s * s;
        nt += t[k] * t[k];
    }
    return nt > 0 ? std::sqrt(res / nt) : std::sqrt(res);
}

// S_kl in the basis {S11,S22,S33,S12,S13,S23} (rows of the barycentric Gram matrix sum to zero)
void s_in_b0(int k, int l, double* v)
{
    for (int i = 0; i < 6; ++i) v[i] = 0;
    if (k > l) std::swap(k, l);
    auto off = [](int a, int b) { return a == 1 ? (b == 2 ? 3 : 4) : 5; };   // (1,2)->3 (1,3)->4 (2,3)->5
    if (k >= 1) {
        if (k == l) v[k - 1] = 1;
        else v[off(k, l)] = 1;
        return;
    }
    if (l >= 1) {   // S_0l = -(S_1l + S_2l + S_3l)
        for (int a = 1; a <= 3; ++a) {
            if (a == l) v[a - 1] -= 1;
            else v[off(std::min(a, l), std::max(a, l))] -= 1;
        }
        return;
    }
    v[0] = v[1] = v[2] = 1;   // S_00 = sum of all S_ab, a,b >= 1
    v[3] = v[4] = v[5] = 2;
}

} // namespace

// ---- building the plan: on the calling thread, or by a background thread (FEELGPU_PATTERN_ASYNC) -------------------------
struct EPlanAsync {
    std::thread th;
    std::atomic<int> state{0};       // 0 building, 1 built, 2 the mesh does not fit the plan (hp.error)
    EPlanHost hp;
    std::unique_ptr<uint16_t[]> pos;  // host copy of the position map, fetched by the builder's own thread
    double seconds = 0;
};

static int eplan_install(feelgpu_pattern_t pat, EPlanHost& hp)
{
    feelgpu_context_t ctx = pat->ctx;
    EPlan* p = new EPlan;
    p->order = hp.order;
    p->n = hp.n;
    p->ncomp = hp.ncomp;
    p->omax = hp.omax;
    p->ncta = (int64_t)hp.ctas.size();
    p->ntasks = hp.ntasks;
    p->niters = hp.niters;
    p->niters_padded = hp.niters_padded;
    p->nvisits = hp.nvisits;
    p->nouts = hp.nouts;
    p->lds_wavefronts = hp.lds_wavefronts;
    p->lds_ideal = hp.lds_ideal;
    cudaError_t err = upload(&p->d_ctas, hp.ctas, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_verts, hp.verts, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_visits, hp.visits, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_chunks, hp.chunks, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_blob, hp.blob, p->bytes);
    if (err == cudaSuccess) err = upload(&p->d_rows, hp.rows, p->bytes);
    p->nrowent = (int64_t)hp.rows.size();
    if (err == cudaSuccess) err = cudaMalloc(&p->d_sched, 2 * sizeof(int));
    if (err == cudaSuccess) err = cudaMemset(p->d_sched, 0, 2 * sizeof(int));
    pat->eplan = p;
    if (err != cudaSuccess) {
        ctx->last_error = std::string("entity plan upload: ") + cudaGetErrorString(err);
        feelgpu_eplan_destroy(pat);
        return FEELGPU_ERR_CUDA;
    }
    return FEELGPU_OK;
}

int feelgpu_eplan_build(feelgpu_pattern_t pat, const int64_t* h_ia, bool async)
{
    feelgpu_context_t ctx = pat->ctx;
    feelgpu_space_t sp = pat->test;
    feelgpu_mesh_t m = sp->mesh;
    if (pat->test != pat->trial) return FEELGPU_OK;   // the plan needs test == trial
    if (!eplan_spec(sp->order)) { pat->eplan_note = "entity plan: order not covered"; return FEELGPU_OK; }
    const int n = sp->n;
    (void)h_ia;                                       // (the pattern keeps its own copy: pat->h_ia)
    EPlanAsync* job = new EPlanAsync;
    if (sp->nDofOwned < pat->nRows) job->hp.nomerge_from = pat->h_ia[sp->nDofOwned];
    const int device = ctx->device, nthreads = feelgpu_builder_threads(ctx);
    auto work = [pat, sp, m, job, n, device, nthreads] {
        const auto t0 = std::chrono::steady_clock::now();
        bool ok = false;
        try {
            // the position map (2 bytes per local matrix entry, 1.6 GB on 8 M P2 elements) comes to the host on the builder's
            // thread and its own stream: with FEELGPU_PATTERN_ASYNC neither the copy nor the page faults of the buffer hold
            // up pattern_create
            const size_t npos = (size_t)m->nElt * n * n;
            job->pos.reset(new uint16_t[std::max<size_t>(npos, 1)]);
            cudaStream_t cs = nullptr;
            cudaError_t err = cudaSetDevice(device);
            if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking);
            if (err == cudaSuccess) err = cudaMemcpyAsync(job->pos.get(), pat->d_pos, npos * sizeof(uint16_t), cudaMemcpyDeviceToHost, cs);
            if (err == cudaSuccess) err = cudaStreamSynchronize(cs);
            if (cs) cudaStreamDestroy(cs);
            if (err != cudaSuccess) job->hp.error = std::string("entity plan: position map to the host: ") + cudaGetErrorString(err);
            else
                ok = eplan_build_host(sp->order, sp->ncomp, EPLAN_OMAX, m->nElt, sp->h_l2g.data(), sp->nd, m->nPts, m->h_xyz.data(),
                                      m->h_conn.data(), pat->nRows, pat->h_ia.data(), job->pos.get(), nthreads, job->hp);
        } catch (const std::exception& e) {
            job->hp.error = std::string("entity plan: ") + e.what();
        }
        job->pos.reset();
        job->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        job->state.store(ok ? 1 : 2, std::memory_order_release);
    };
    pat->eplan_async = job;
    if (async) {
        job->th = std::thread(work);
        return FEELGPU_OK;
    }
    work();
    return feelgpu_eplan_poll(pat, true);
}

// installs the plan of a finished background build (wait: block until it is finished)
int feelgpu_eplan_poll(feelgpu_pattern_t pat, bool wait)
{
    EPlanAsync* job = pat->eplan_async;
    if (!job) return FEELGPU_OK;
    if (!wait && job->state.load(std::memory_order_acquire) == 0) return FEELGPU_OK;
    if (job->th.joinable()) job->th.join();
    pat->eplan_async = nullptr;
    int rc = FEELGPU_OK;
    pat->eplan_build_seconds = job->seconds;
    if (job->state.load() == 1) {
        cudaSetDevice(pat->ctx->device);
        rc = eplan_install(pat, job->hp);
    } else
        pat->eplan_note = job->hp.error;   // not an error: another path stays available
    delete job;
    return rc;
}

int feelgpu_eplan_stats(feelgpu_pattern_t pat, int64_t stats[10])
{
    for (int k = 0; k < 10; ++k) stats[k] = 0;
    const EPlan* p = pat->eplan;
    if (!p) return FEELGPU_OK;
    stats[0] = 1;
    stats[1] = p->ncta;
    stats[2] = p->ntasks;
    stats[3] = p->niters;
    stats[4] = p->niters_padded;
    stats[5] = p->nvisits;
    stats[6] = p->nouts;
    stats[7] = p->bytes;
    stats[8] = (int64_t)p->lds_wavefronts;
    stats[9] = (int64_t)p->lds_ideal;
    return FEELGPU_OK;
}

void feelgpu_eplan_destroy(feelgpu_pattern_t pat)
{
    if (EPlanAsync* job = pat->eplan_async) {      // a build still running: let it finish, drop the result
        if (job->th.joinable()) job->th.join();
        delete job;
        pat->eplan_async = nullptr;
    }
    EPlan* p = pat->eplan;
    if (!p) return;
    cudaFree(p->d_ctas);
    cudaFree(p->d_verts);
    cudaFree(p->d_visits);
    cudaFree(p->d_chunks);
    cudaFree(p->d_blob);
    cudaFree(p->d_rows);
    cudaFree(p->d_rows_local);
    cudaFree(p->d_sched);
    delete p;
    pat->eplan = nullptr;
}

// coefficients of every (type, output) from the caller's reference tables; false if the basis does not have the
// entity-local structure the plan assumes (the residual of the projection onto the loaded S entries is not zero)
static bool eplan_coefficients(const EPlanSpec& S, const feelgpu_tables& t, int form, ECoef* out, std::string* why)
{
    std::memset(out, 0, sizeof(ECoef));
    const int nq = t.nq;
    for (int ty = 0; ty < S.ntypes; ++ty) {
        const EPlanType& T = S.type[ty];
        const int i0 = T.rowdof;
        for (int o = 0; o < T.nout; ++o) {
            const int j = T.outdof[o];
            if (form == 0) {
                double mm = 0;
                for (int q = 0; q < nq; ++q) mm += t.w[q] * (t.phi[i0 * nq + q] * t.phi[j * nq + q]);
                out->mass[ty][o] = mm;
                continue;
            }
            double M[3][3];
            for (int b = 0; b < 3; ++b)
                for (int d = 0; d < 3; ++d) {
                    double v = 0;
                    for (int q = 0; q < nq; ++q) v += t.w[q] * (t.dphi[(i0 * nq + q) * 3 + b] * t.dphi[(j * nq + q) * 3 + d]);
                    M[b][d] = v;
                }
            // A = sum_bd G'_bd M_bd with G'_bd = 4 S_{b+1,d+1}  (grad xi_b = 2 grad lambda_{b+1})
            const double tv[6] = {4 * M[0][0], 4 * M[1][1], 4 * M[2][2], 4 * (M[0][1] + M[1][0]), 4 * (M[0][2] + M[2][0]),
                                  4 * (M[1][2] + M[2][1])};
            double u[EPLAN_MAXLD][6], x[EPLAN_MAXLD];
            for (int k = 0; k < T.nld; ++k) s_in_b0(T.ld[k][0], T.ld[k][1], u[k]);
            const double res = lsq6(T.nld, u, tv, x);
            if (!(res < 1e-11)) {
                *why = "entity plan: the reference tables are not those of an entity-local Lagrange basis";
                return false;
            }
            for (int k = 0; k < T.nld; ++k) out->coef[ty][o][k] = x[k];
            if (T.phaseA)   // phase A applies one coefficient to the sum of the loaded entries
                for (int k = 1; k < T.nld; ++k)
                    if (std::fabs(x[k] - x[0]) > 1e-11 * std::fabs(x[0])) {
                        *why = "entity plan: the (edge, opposite edge) entry is not symmetric in the loaded S entries";
                        return false;
                    }
        }
    }
    return true;
}

bool feelgpu_eplan_supports(feelgpu_pattern_t pat, const feelgpu_form2_desc* desc)
{
    if (!pat->eplan) return false;
    if (desc->kind != FEELGPU_FORM2_MASS && desc->kind != FEELGPU_FORM2_STIFFNESS) return false;
    if (desc->coeff.kind != FEELGPU_COEFF_CONST) return false;
    const EPlanSpec* S = eplan_spec(pat->eplan->order);
    if (!S || desc->tables.n != S->n) return false;
    if (desc->kind == FEELGPU_FORM2_STIFFNESS && !desc->tables.dphi) return false;
    ECoef tmp;
    std::string why;
    return eplan_coefficients(*S, desc->tables, desc->kind == FEELGPU_FORM2_STIFFNESS ? 1 : 0, &tmp, &why);
}

int feelgpu_launch_form2_eplan(feelgpu_pattern_t pat, const FormParams& hp, const feelgpu_form2_desc* desc,
                               int accumulate, double* a, cudaStream_t s)
{
    feelgpu_context_t ctx = pat->ctx;
    EPlan* p = pat->eplan;
    feelgpu_mesh_t m = pat->test->mesh;
    const int form = desc->kind == FEELGPU_FORM2_STIFFNESS ? 1 : 0;
    ECoef tab;
    std::string why;
    if (!eplan_coefficients(*eplan_spec(p->order), desc->tables, form, &tab, &why)) {
        ctx->last_error = why;
        return FEELGPU_ERR_UNSUPPORTED;
    }
    // stream-ordered upload of the coefficients (one context drives one stream at a time, see feelgpu.h)
    FEELGPU_CHECK_CUDA(ctx, cudaMemcpyToSymbolAsync(c_ecoef, &tab, sizeof(ECoef), 0, cudaMemcpyHostToDevice, s));
    ERemote rem;
    rem.n = p->nremote;
    rem.accumulate = accumulate && !p->remote_fresh;   // a landing buffer starts every epoch empty, whatever it holds
    p->remote_fresh = false;
    for (int k = 0; k < EPLAN_MAXREMOTE; ++k) rem.ptr[k] = k < p->nremote ? p->remote_ptr[p->remote_epoch & 1][k] : nullptr;
    const size_t sh = eplan_smem_bytes(p, form);
    // persistent CTAs: EPLAN_MINB per SM, each runs every gridDim.x-th patch
    // stiffness: when the S array (behind the row block and the chunk headers) starts at shared address 0x10000 its load
    // addresses are composed from bit fields of the unit word
    const int sbase = smem_base(ctx, s);
    const bool al = EPLAN_VS == 512 && sbase >= 0 &&
                    sbase + (int64_t)p->omax * 8 + EPLAN_HDRS * (int64_t)sizeof(uint2) == 0x10000 && !getenv("FEELGPU_EPLAN_NO_ALIGNED_S");
    const int grid = (int)std::min<int64_t>(p->ncta, (int64_t)ctx->sm_count * EPLAN_MINB), threads = EPLAN_WARPS * 32;
    if (grid == 0) return FEELGPU_OK;
#define ELAUNCH1(O, F, C, A)                                                                                         \
    do {                                                                                                            \
        FEELGPU_CHECK_CUDA(ctx, cudaFuncSetAttribute(k_eplan<O, F, C, A>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh)); \
        k_eplan<O, F, C, A><<<grid, threads, sh, s>>>(p->d_ctas, (int)p->ncta, p->d_sched, p->d_verts, p->d_visits,                              \
                                                   reinterpret_cast<const uint2*>(p->d_chunks), p->d_blob,          \
                                                   reinterpret_cast<const uint4*>(p->d_rows), m->d_xyz, a, accumulate, \
                                                   EPLAN_VS, p->omax, hp.cval[0], rem);                                  \
    } while (0)
#define ELAUNCH(O, F)                         \
    do {                                      \
        if (p->ncomp == 1) { if (F == 1 && al) ELAUNCH1(O, F, 1, true); else ELAUNCH1(O, F, 1, false); } \
        else { if (F == 1 && al) ELAUNCH1(O, F, 3, true); else ELAUNCH1(O, F, 3, false); } \
    } while (0)
    switch (p->order * 2 + form) {
    case 2: ELAUNCH(1, 0); break;
    case 3: ELAUNCH(1, 1); break;
    case 4: ELAUNCH(2, 0); break;
    case 5: ELAUNCH(2, 1); break;
    default: ctx->last_error = "entity plan: unsupported order"; return FEELGPU_ERR_UNSUPPORTED;
    }
#undef ELAUNCH
#undef ELAUNCH1
    ctx->launches++;
    FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
    return FEELGPU_OK;
}

// ---- ghost rows into the owners' landing buffers --------------------------------------------------------------------------
namespace {
// every row-table entry of a ghost row (base at or behind tail0: one entry per node row, never merged) gets
// EPLAN_REMOTE_FLAG | neighbour << 56 | offset; the dof row is found by its CSR offset
__global__ void k_eplan_patch_rows(int64_t n, const EPlanRow* __restrict__ local, EPlanRow* __restrict__ rows, int64_t tail0,
                                   const int64_t* __restrict__ ia, int64_t nOwned, int64_t nRows,
                                   const int32_t* __restrict__ ghost_nb, const int64_t* __restrict__ ghost_off, int* __restrict__ flag)
{
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
        EPlanRow r = local[k];
        if (ghost_nb && r.base >= tail0) {
            int64_t lo = nOwned, hi = nRows;             // first row with ia[row] >= base (ghost rows are never empty)
            while (lo < hi) {
                const int64_t mid = (lo + hi) >> 1;
                if (ia[mid] < r.base) lo = mid + 1;
                else hi = mid;
            }
            if (lo >= nRows || ia[lo] != r.base) atomicExch(flag, 1);
            else r.base = EPLAN_REMOTE_FLAG | ((int64_t)ghost_nb[lo - nOwned] << 56) | ghost_off[lo - nOwned];
        }
        rows[k] = r;
    }
}
} // namespace

int feelgpu_eplan_set_remote(feelgpu_pattern_t pat, int n, const int32_t* h_ghost_nb, const int64_t* h_ghost_off,
                             double* const* ptr0, double* const* ptr1)
{
    feelgpu_context_t ctx = pat->ctx;
    EPlan* p = pat->eplan;
    if (!p) {
        if (n == 0) return FEELGPU_OK;
        ctx->last_error = "set_remote: the pattern has no entity plan (only that kernel writes to peer memory)";
        return FEELGPU_ERR_UNSUPPORTED;
    }
    if (n < 0 || n > EPLAN_MAXREMOTE) { ctx->last_error = "set_remote: too many owners"; return FEELGPU_ERR_ARG; }
    cudaStream_t s = ctx->stream;
    const int64_t nO = pat->test->nDofOwned, nG = pat->nRows - nO;
    if (!p->d_rows_local) {     // keep the table as built
        FEELGPU_CHECK_CUDA(ctx, cudaMalloc(&p->d_rows_local, std::max<int64_t>(p->nrowent, 1) * sizeof(EPlanRow)));
        FEELGPU_CHECK_CUDA(ctx, cudaMemcpyAsync(p->d_rows_local, p->d_rows, p->nrowent * sizeof(EPlanRow), cudaMemcpyDeviceToDevice, s));
    }
    int32_t* d_nb = nullptr;
    int64_t* d_off = nullptr;
    if (n > 0 && nG > 0) {
        FEELGPU_CHECK_CUDA(ctx, cudaMalloc(&d_nb, nG * sizeof(int32_t)));
        FEELGPU_CHECK_CUDA(ctx, cudaMalloc(&d_off, nG * sizeof(int64_t)));
        FEELGPU_CHECK_CUDA(ctx, cudaMemcpyAsync(d_nb, h_ghost_nb, nG * sizeof(int32_t), cudaMemcpyHostToDevice, s));
        FEELGPU_CHECK_CUDA(ctx, cudaMemcpyAsync(d_off, h_ghost_off, nG * sizeof(int64_t), cudaMemcpyHostToDevice, s));
    }
    int flag = 0;
    if (p->nrowent > 0) {
        FEELGPU_CHECK_CUDA(ctx, cudaMemsetAsync(ctx->d_flag, 0, sizeof(int), s));
        const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((p->nrowent + 255) / 256, (int64_t)ctx->sm_count * 8));
        k_eplan_patch_rows<<<grid, 256, 0, s>>>(p->nrowent, p->d_rows_local, p->d_rows, pat->h_ia[nO], pat->d_ia, nO, pat->nRows,
                                                d_nb, d_off, ctx->d_flag);
        ctx->launches++;
        FEELGPU_CHECK_CUDA(ctx, cudaGetLastError());
        FEELGPU_CHECK_CUDA(ctx, cudaMemcpyAsync(&flag, ctx->d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
    }
    FEELGPU_CHECK_CUDA(ctx, cudaStreamSynchronize(s));
    cudaFree(d_nb);
    cudaFree(d_off);
    if (flag) { ctx->last_error = "set_remote: a ghost row of the plan does not start at a CSR row"; return FEELGPU_ERR_ARG; }
    p->nremote = (nG > 0) ? n : 0;
    p->remote_fresh = true;
    p->remote_epoch = 0;
    for (int k = 0; k < EPLAN_MAXREMOTE; ++k) {
        p->remote_ptr[0][k] = k < n && ptr0 ? ptr0[k] : nullptr;
        p->remote_ptr[1][k] = k < n && ptr1 ? ptr1[k] : nullptr;
    }
    return FEELGPU_OK;
}

void feelgpu_eplan_remote_epoch(feelgpu_pattern_t pat, int epoch)
{
    if (!pat->eplan) return;
    pat->eplan->remote_epoch = epoch;
    pat->eplan->remote_fresh = true;
}

void feelgpu_eplan_remote_fresh(feelgpu_pattern_t pat)
{
    if (pat->eplan) pat->eplan->remote_fresh = true;
}

// ---- host-side view of the plan (include/feelgpu.h) ----------------------------------------------------------------
extern "C" {

int feelgpu_eplan_host_build(int order, int ncomp, int64_t nElt, const int32_t* l2g, int64_t nPts, const double* xyz,
                             const int32_t* conn, int64_t nRows, const int64_t* ia, const uint16_t* pos, void** plan,
                             int64_t sizes[8])
{
    if (!l2g || !xyz || !conn || !ia || !pos || !plan || !sizes) return FEELGPU_ERR_ARG;
    const EPlanSpec* S = eplan_spec(order);
    if (!S) return FEELGPU_ERR_UNSUPPORTED;
    EPlanHost* hp = new EPlanHost;
    *plan = hp;
    const bool ok = eplan_build_host(order, ncomp, EPLAN_OMAX, nElt, l2g, S->n * ncomp, nPts, xyz, conn, nRows, ia, pos, 0, *hp);
    sizes[0] = (int64_t)hp->ctas.size();
    sizes[1] = (int64_t)hp->verts.size();
    sizes[2] = (int64_t)hp->visits.size();
    sizes[3] = (int64_t)hp->chunks.size();
    sizes[4] = (int64_t)hp->blob.size();
    sizes[5] = (int64_t)hp->rows.size();
    sizes[6] = hp->omax;
    sizes[7] = EPLAN_VS;
    return ok ? FEELGPU_OK : FEELGPU_ERR_UNSUPPORTED;
}

int feelgpu_eplan_host_copy(void* plan, void* ctas, int32_t* verts, uint32_t* visits, void* chunks, uint32_t* blob,
                            void* rows, double stats[8])
{
    if (!plan) return FEELGPU_ERR_ARG;
    const EPlanHost* hp = static_cast<const EPlanHost*>(plan);
    if (ctas) std::memcpy(ctas, hp->ctas.data(), hp->ctas.size() * sizeof(EPlanCta));
    if (verts) std::memcpy(verts, hp->verts.data(), hp->verts.size() * sizeof(int32_t));
    if (visits) std::memcpy(visits, hp->visits.data(), hp->visits.size() * sizeof(uint32_t));
    if (chunks) std::memcpy(chunks, hp->chunks.data(), hp->chunks.size() * sizeof(EPlanChunk));
    if (blob) std::memcpy(blob, hp->blob.data(), hp->blob.size() * sizeof(uint32_t));
    if (rows) std::memcpy(rows, hp->rows.data(), hp->rows.size() * sizeof(EPlanRow));
    if (stats) {
        stats[0] = (double)hp->ntasks;
        stats[1] = (double)hp->niters;
        stats[2] = (double)hp->niters_padded;
        stats[3] = (double)hp->nvisits;
        stats[4] = (double)hp->nouts;
        stats[5] = hp->lds_wavefronts;
        stats[6] = hp->lds_ideal;
        stats[7] = hp->a1_passes_before > 0 ? hp->a1_passes / hp->a1_passes_before : 0;
    }
    return FEELGPU_OK;
}

const char* feelgpu_eplan_host_error(void* plan) { return plan ? static_cast<const EPlanHost*>(plan)->error.c_str() : ""; }

int feelgpu_eplan_host_free(void* plan)
{
    delete static_cast<EPlanHost*>(plan);
    return FEELGPU_OK;
}

int feelgpu_eplan_host_coefficients(int order, const feelgpu_tables* tables, int form, double* coef, double* mass)
{
    const EPlanSpec* S = eplan_spec(order);
    if (!S || !tables || !coef || !mass || tables->n != S->n) return FEELGPU_ERR_ARG;
    ECoef tab;
    std::string why;
    if (!eplan_coefficients(*S, *tables, form, &tab, &why)) return FEELGPU_ERR_UNSUPPORTED;
    std::memcpy(coef, tab.coef, sizeof(tab.coef));
    std::memcpy(mass, tab.mass, sizeof(tab.mass));
    return FEELGPU_OK;
}

} // extern "C"
