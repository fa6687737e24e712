// blockpool.cuh — block-level ray pool: the traversal core of every trace kernel (replaces the per-warp lane pool of
// traverse.cuh, kept for A/B with -DQB_BLOCKPOOL=0).
//
// The per-ray work is the same flat state machine as traverse.cuh — Scene::TraceNodeNormal / TraceNodeShadow
// (src/scene/scene.cpp:35-74): object k of the pre-order object list -> Node::ToNodeCoords chain -> Sphere | Plane | TriObj
// box + TriObj::TraceBVHNode (src/objects/objects.cpp:324-420, near child first, far child deferred) — with every visit,
// triangle test and comparison in the reference's order, so hit/miss, primitive id and t stay bit-identical. What changes is
// WHERE a ray lives and WHO advances it:
//
//   * A ray's state (33 words + its stack of deferred far children) lives in SHARED memory, in slot (row, column) of a pool
//     of rows x 32 slots per block; column = lane index, so any warp's lane l can work on any of the `rows` rays of column l
//     and every state access is bank-conflict free.
//   * Each slot is in one of four phases — FREE (no ray), INNER (current node is an internal BVH node), LEAF (a leaf),
//     NEXT (between objects) — published as one availability bit per slot in avail[phase][column].
//   * A warp votes for the phase with the most columns that have work, every lane claims a slot of that phase in its own
//     column (one shared-memory atomic), runs the phase on it and releases it into its new phase. In the INNER phase a lane
//     whose ray leaves the phase claims the next inner ray of its column at once, so the warp stays full while the block
//     holds enough rays of that kind.
//
// The lane pool kept one ray per lane in registers and ran the phase most lanes were waiting for: with rays spread over
// three kinds of work, 13-19 of 32 lanes were active per instruction (ncu, profiles/ncu_trace_r2a_summary.txt). Here the
// lanes of a warp always execute the same phase on whatever rays of that phase the block has.
// Must be compiled with --fmad=false like trace.cuh.
#pragma once
#include <cstdio>

#include "traverse.cuh"

namespace qb {

#ifndef QB_BP_WARPS
#define QB_BP_WARPS 8                 // warps per block; two blocks per SM (shadow(d) runs beside trace(d + 1))
#endif
#define QB_BP_THREADS (QB_BP_WARPS * 32)
#ifndef QB_BP_BLOCKS
#define QB_BP_BLOCKS 2                // resident blocks per SM
#endif
#ifndef QB_BP_SMEM
#define QB_BP_SMEM (100 * 1024)       // dynamic shared memory per block: pool state + stacks
#endif
#ifndef QB_BP_INNER_REPS
#define QB_BP_INNER_REPS 16           // node visits per vote (lanes refill from their column between visits)
#endif
#ifndef QB_BP_REFILL
#define QB_BP_REFILL 8                // idle lanes that trigger a refill of the INNER loop
#endif
#ifndef QB_BP_MIN_LANES
#define QB_BP_MIN_LANES 12            // the INNER loop ends when fewer lanes than this still hold a ray
#endif
#ifndef QB_BP_W_LEAF
#define QB_BP_W_LEAF 4                // vote weights in quarters (INNER = 4)
#endif
#ifndef QB_BP_W_NEXT
#define QB_BP_W_NEXT 4
#endif
#ifndef QB_BP_W_FETCH
#define QB_BP_W_FETCH 4
#endif

enum { BP_FREE = 0, BP_INNER = 1, BP_LEAF = 2, BP_NEXT = 3 };
// words of a slot's state
enum {
  F_IDX = 0,
  F_C0PX, F_C0PY, F_C0PZ, F_C0DX, F_C0DY, F_C0DZ,   // the ray in the root node's space
  F_RPX, F_RPY, F_RPZ, F_RDX, F_RDY, F_RDZ,         // ... in the current object's space
  F_DCX, F_DCY, F_DCZ,                              // 1 / dir of the current object's ray (objects.cpp:333)
  F_Z, F_CUR, F_SP, F_K, F_FLAGS,                   // flags: bit 0-2 degenerate axes x y z, bits 3-4 cull_state
  F_MEXIT, F_ZENTRY, F_ZSANE,
  F_OBJSLOT, F_UVSLOT, F_UVT,
  F_HK, F_HSLOT, F_HTEXK, F_HTEXSLOT, F_HTEXT, F_HLASTTRI,
#ifdef QB_BP_DEBUG
  F_DBG_VISITS, F_DBG_NEXTS,
#endif
  F_COUNT
};
#define QB_BP_HDR_WORDS 160  // avail[4][32] + control words

struct BlockPool {
  uint32_t *avail;   // [4][32]
  uint32_t *ctrl;    // [0] ray source exhausted, [1] rays alive in the pool
  uint32_t *coop;    // this warp's leaf scratch (QB_COOP_SLOTS words)
  uint32_t *state;   // [F_COUNT][rows][32]
  uint32_t *stack;   // [entries][rows][32]
  int rows;
};

// rows of the pool that fit QB_BP_SMEM with `entries` stack entries per ray (host and device)
__host__ __device__ inline int bp_rows(int entries)
{
  const int fixed = (QB_BP_HDR_WORDS + QB_BP_WARPS * QB_COOP_SLOTS) * 4;
  int rows = (QB_BP_SMEM - fixed) / (32 * 4 * (F_COUNT + entries));
  return rows > 32 ? 32 : rows;
}
__host__ __device__ inline size_t bp_smem_bytes(int entries)
{
  return (size_t) (QB_BP_HDR_WORDS + QB_BP_WARPS * QB_COOP_SLOTS) * 4 + (size_t) bp_rows(entries) * 32 * 4 * (F_COUNT + entries);
}

__device__ __forceinline__ uint32_t bp_ld(const uint32_t *p) { return *(const volatile uint32_t *) p; }

// claim one slot of `phase` in this lane's column; returns the row or -1. m = the availability word read for the vote.
__device__ __forceinline__ int bp_claim(const BlockPool &P, int phase, int lane, uint32_t m, int start)
{
  if (m == 0u) return -1;
  const uint32_t hi = m & ~((1u << start) - 1u);
  const int row = __ffs(hi ? hi : m) - 1;  // different warps start at different rows: fewer lost races
  const uint32_t old = atomicAnd(&P.avail[phase * 32 + lane], ~(1u << row));
  return ((old >> row) & 1u) ? row : -1;
}
__device__ __forceinline__ void bp_release(const BlockPool &P, int phase, int lane, int row)
{
  __threadfence_block();  // the slot's state is visible before its availability bit
  atomicOr(&P.avail[phase * 32 + lane], 1u << row);
}
__device__ __forceinline__ int bp_phase_of(uint32_t cur) { return cur == QB_REF_EMPTY ? BP_NEXT : ((cur & QB_LEAF_BIT) ? BP_LEAF : BP_INNER); }

#define ST(f) P.state[((f) * P.rows + row) * 32 + lane]
#define STF(f) __uint_as_float(ST(f))
#define STK(e) P.stack[((e) * P.rows + row) * 32 + lane]

// IO concept: as traverse.cuh (count, cursor, map, load, store, store_any, defer, exact); every call names the ray by index,
// so an IO object must not keep per-ray state between calls (a ray is advanced by different threads over its life).
template <bool ANY, class IO>
__device__ __forceinline__ void traverse_block(const DScene &S, IO &io, uint32_t *sm, int entries, const SceneTab &tab, TraceCounters &cnt)
{
  TabRef T = tab_ref(tab);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t FULL = 0xFFFFFFFFu;
  BlockPool P;
  P.avail = sm;
  P.ctrl = sm + 128;
  P.coop = sm + QB_BP_HDR_WORDS + warp * QB_COOP_SLOTS;
  P.state = sm + QB_BP_HDR_WORDS + QB_BP_WARPS * QB_COOP_SLOTS;
  P.rows = bp_rows(entries);
  P.stack = P.state + F_COUNT * P.rows * 32;
  const uint32_t n = io.count();
  if (threadIdx.x < 128) sm[threadIdx.x] = threadIdx.x < 32 ? (P.rows >= 32 ? FULL : (1u << P.rows) - 1u) : 0u;
  if (threadIdx.x >= 128 && threadIdx.x < QB_BP_HDR_WORDS) sm[threadIdx.x] = 0u;
  __syncthreads();
  const int start = (warp * P.rows) / QB_BP_WARPS;

#ifdef QB_BP_DEBUG
  unsigned long long iters = 0, it_sleep = 0, it_inner = 0, it_leaf = 0, it_next = 0, it_fetch = 0, claims = 0;
#endif
  for (;;) {
#ifdef QB_BP_DEBUG
    if (++iters > 4000000ull) {
      {
        const unsigned long long cl = __reduce_add_sync(FULL, (unsigned) claims);
        if (lane == 0) printf("  w%d iterations: sleep %llu inner %llu leaf %llu next %llu fetch %llu claims %llu\n", warp, it_sleep, it_inner, it_leaf, it_next, it_fetch, cl);
      }
      if (lane == 0)
        printf("blockpool stuck: block %d warp %d n %u cursor %u exhausted %u live %u avail F %08x I %08x L %08x N %08x\n", blockIdx.x, warp, n,
               *io.cursor(), bp_ld(&P.ctrl[0]), bp_ld(&P.ctrl[1]), bp_ld(&P.avail[0]), bp_ld(&P.avail[32]), bp_ld(&P.avail[64]), bp_ld(&P.avail[96]));
      {
        const uint32_t freem = bp_ld(&P.avail[lane]);
        for (int row = 0; row < P.rows; ++row)
          if (!((freem >> row) & 1u))
            printf("  w%d busy slot col %d row %d: I %x L %x N %x idx %u k %u cur %08x sp %u flags %x z %g visits %u nexts %u zentry %g zsane %g mexit %g\n", warp, lane, row, (bp_ld(&P.avail[32 + lane]) >> row) & 1u, (bp_ld(&P.avail[64 + lane]) >> row) & 1u, (bp_ld(&P.avail[96 + lane]) >> row) & 1u, ST(F_IDX), ST(F_K),
                   ST(F_CUR), ST(F_SP), ST(F_FLAGS), STF(F_Z), ST(F_DBG_VISITS), ST(F_DBG_NEXTS), STF(F_ZENTRY), STF(F_ZSANE), STF(F_MEXIT));
      }
      break;
    }
#endif
    const bool exhausted = bp_ld(&P.ctrl[0]) != 0u;
    const uint32_t aF = exhausted ? 0u : bp_ld(&P.avail[BP_FREE * 32 + lane]);
    const uint32_t aI = bp_ld(&P.avail[BP_INNER * 32 + lane]), aL = bp_ld(&P.avail[BP_LEAF * 32 + lane]), aN = bp_ld(&P.avail[BP_NEXT * 32 + lane]);
    const int n_fetch = __popc(__ballot_sync(FULL, aF != 0u)), n_inner = __popc(__ballot_sync(FULL, aI != 0u));
    const int n_leaf = __popc(__ballot_sync(FULL, aL != 0u)), n_next = __popc(__ballot_sync(FULL, aN != 0u));
    if ((n_fetch | n_inner | n_leaf | n_next) == 0) {
      if (exhausted && bp_ld(&P.ctrl[1]) == 0u) break;  // no ray left anywhere in the block
      __nanosleep(40);  // every ray of the block is in another warp's hands
#ifdef QB_BP_DEBUG
      ++it_sleep;
#endif
      continue;
    }
    const int w_inner = 4 * n_inner, w_leaf = QB_BP_W_LEAF * n_leaf, w_next = QB_BP_W_NEXT * n_next, w_fetch = QB_BP_W_FETCH * n_fetch;

    if (w_inner >= w_leaf && w_inner >= w_next && w_inner >= w_fetch) {
      // ---- INNER: node visits (both children's boxes); the near child becomes current, the far child is deferred -------
#ifdef QB_BP_DEBUG
      ++it_inner;
#endif
      int row = -1;
      const bool alone = (n_leaf | n_next | n_fetch) == 0;
      Ray3 r;
      V3 drcp = v3(0.f);
      float z = 0.f, mesh_exit = 0.f;
      uint32_t cur = QB_REF_EMPTY, flags = 0u;
      int sp = 0;
      bool done = false;  // this lane's ray has left the phase (a leaf, or its mesh is finished): handed on at the next refill
#pragma unroll 1
      for (int rep = 0; rep < QB_BP_INNER_REPS; ++rep) {
        // Hand-over and refill are batched: with a lane or two at a time they cost as many issue slots as a node visit
        // (ncu: 20 % of the kernel's instructions at 5-7 of 32 lanes). Lanes whose ray has left wait until QB_BP_REFILL
        // lanes are idle, then all of them pass their ray on and take the next inner ray of their column.
        const int n_idle = __popc(__ballot_sync(FULL, row < 0 || done));
        if (rep == 0 || n_idle >= QB_BP_REFILL) {
          if (done) {
            ST(F_CUR) = cur;
            ST(F_SP) = (uint32_t) sp;
            ST(F_FLAGS) = flags;
            bp_release(P, bp_phase_of(cur), lane, row);
            row = -1;
            done = false;
          }
          if (row < 0) {
            const uint32_t m = rep == 0 ? aI : bp_ld(&P.avail[BP_INNER * 32 + lane]);
            row = bp_claim(P, BP_INNER, lane, m, start);
#ifdef QB_BP_DEBUG
            if (row >= 0) ++claims;
#endif
            if (row >= 0) {
              __threadfence_block();
              r.p = v3(STF(F_RPX), STF(F_RPY), STF(F_RPZ));
              drcp = v3(STF(F_DCX), STF(F_DCY), STF(F_DCZ));
              z = STF(F_Z);
              cur = ST(F_CUR);
              sp = (int) ST(F_SP);
              flags = ST(F_FLAGS);
              if (flags & 7u) {
                r.d = v3(STF(F_RDX), STF(F_RDY), STF(F_RDZ));
                mesh_exit = STF(F_MEXIT);
              }
            }
          }
        }
        const bool go = row >= 0 && !done;
        if (go) {
          const DBvhNode *np = &S.bvh[cur];
          const float4 na = __ldg(&np->a), nb = __ldg(&np->b), nc = __ldg(&np->c);
          const uint4 ndr = __ldg(&np->d);
          QB_CNT(nodes);
#ifdef QB_BP_DEBUG
          ST(F_DBG_VISITS) = ST(F_DBG_VISITS) + 1u;
#endif
          const float p00x = -(r.p.x - na.x) * drcp.x, p00y = -(r.p.y - na.y) * drcp.y, p00z = -(r.p.z - na.z) * drcp.z;
          const float p01x = -(r.p.x - na.w) * drcp.x, p01y = -(r.p.y - nb.x) * drcp.y, p01z = -(r.p.z - nb.y) * drcp.z;
          const float p10x = -(r.p.x - nb.z) * drcp.x, p10y = -(r.p.y - nb.w) * drcp.y, p10z = -(r.p.z - nc.x) * drcp.z;
          const float p11x = -(r.p.x - nc.y) * drcp.x, p11y = -(r.p.y - nc.z) * drcp.y, p11z = -(r.p.z - nc.w) * drcp.z;
          const float t_max = z;
          float entry0, entry1, exit0, exit1;
          bool h0, h1;
          if (!(flags & 7u)) {
            // no degenerate axis: every p is finite or +-inf, never NaN, so the reference's MIN/MAX macros
            // (x < y ? x : y) and fminf/fmaxf pick the same value up to the sign of a zero, which no comparison sees
            entry0 = fmaxf(fminf(p00x, p01x), fmaxf(fminf(p00y, p01y), fminf(p00z, p01z)));
            entry1 = fmaxf(fminf(p10x, p11x), fmaxf(fminf(p10y, p11y), fminf(p10z, p11z)));
            exit0 = fminf(fmaxf(p00x, p01x), fminf(fmaxf(p00y, p01y), fmaxf(p00z, p01z)));
            exit1 = fminf(fmaxf(p10x, p11x), fminf(fmaxf(p10y, p11y), fmaxf(p10z, p11z)));
            h0 = (entry0 < t_max && entry0 < exit0);
            h1 = (entry1 < t_max && entry1 < exit1);
          } else {
            // degenerate axis (|d| < 1e-7): the slab test ignores it; boxes the ray cannot reach along it are dropped and the
            // (ray, mesh) pair is checked exactly afterwards — see traverse.cuh for the argument, which applies unchanged
            const bool gx = flags & 1u, gy = flags & 2u, gz = flags & 4u;
            const float t00x = gx ? -QB_BIGFLOAT : qmin(p00x, p01x), t01x = gx ? QB_BIGFLOAT : qmax(p00x, p01x);
            const float t10x = gx ? -QB_BIGFLOAT : qmin(p10x, p11x), t11x = gx ? QB_BIGFLOAT : qmax(p10x, p11x);
            const float t00y = gy ? -QB_BIGFLOAT : qmin(p00y, p01y), t01y = gy ? QB_BIGFLOAT : qmax(p00y, p01y);
            const float t10y = gy ? -QB_BIGFLOAT : qmin(p10y, p11y), t11y = gy ? QB_BIGFLOAT : qmax(p10y, p11y);
            const float t00z = gz ? -QB_BIGFLOAT : qmin(p00z, p01z), t01z = gz ? QB_BIGFLOAT : qmax(p00z, p01z);
            const float t10z = gz ? -QB_BIGFLOAT : qmin(p10z, p11z), t11z = gz ? QB_BIGFLOAT : qmax(p10z, p11z);
            entry0 = qmax(t00x, qmax(t00y, t00z)); entry1 = qmax(t10x, qmax(t10y, t10z));
            exit0 = qmin(t01x, qmin(t01y, t01z)); exit1 = qmin(t11x, qmin(t11y, t11z));
            h0 = (entry0 < t_max && entry0 < exit0);
            h1 = (entry1 < t_max && entry1 < exit1);
            const float t_far = qmin(t_max, mesh_exit);
            const bool h0r = h0, h1r = h1;
            if (((flags >> 3) & 3u) != 2u) {
              if (gx) {
                const float mg = 2.f * qabs(r.d.x) * t_far;
                const float m0 = mg + 1e-4f * qmax(qabs(r.p.x), qmax(qabs(na.x), qabs(na.w)));
                const float m1 = mg + 1e-4f * qmax(qabs(r.p.x), qmax(qabs(nb.z), qabs(nc.y)));
                h0 = h0 && !(r.p.x < na.x - m0 || r.p.x > na.w + m0);
                h1 = h1 && !(r.p.x < nb.z - m1 || r.p.x > nc.y + m1);
              }
              if (gy) {
                const float mg = 2.f * qabs(r.d.y) * t_far;
                const float m0 = mg + 1e-4f * qmax(qabs(r.p.y), qmax(qabs(na.y), qabs(nb.x)));
                const float m1 = mg + 1e-4f * qmax(qabs(r.p.y), qmax(qabs(nb.w), qabs(nc.z)));
                h0 = h0 && !(r.p.y < na.y - m0 || r.p.y > nb.x + m0);
                h1 = h1 && !(r.p.y < nb.w - m1 || r.p.y > nc.z + m1);
              }
              if (gz) {
                const float mg = 2.f * qabs(r.d.z) * t_far;
                const float m0 = mg + 1e-4f * qmax(qabs(r.p.z), qmax(qabs(na.z), qabs(nb.y)));
                const float m1 = mg + 1e-4f * qmax(qabs(r.p.z), qmax(qabs(nc.x), qabs(nc.w)));
                h0 = h0 && !(r.p.z < na.z - m0 || r.p.z > nb.y + m0);
                h1 = h1 && !(r.p.z < nc.x - m1 || r.p.z > nc.w + m1);
              }
              if ((h0r && !h0) || (h1r && !h1)) flags = (flags & ~0x18u) | (1u << 3);  // cull_state = 1: something was dropped
            }
          }
          {
            const bool both = h0 && h1, none = !(h0 || h1);
            const bool first0 = entry0 < entry1;  // the reference pushes the other child first, then this one
            const bool take0 = both ? first0 : h0;
            const uint32_t near_ref = take0 ? ndr.x : ndr.y, far_ref = take0 ? ndr.y : ndr.x;
            if (both) STK(sp) = far_ref;
            sp += both ? 1 : 0;
            const bool pop = none && sp > 0;
            sp -= pop ? 1 : 0;
            uint32_t popped = QB_REF_EMPTY;
            if (pop) popped = STK(sp);
            cur = none ? popped : near_ref;
          }
          done = cur == QB_REF_EMPTY || (cur & QB_LEAF_BIT) != 0u;  // the ray leaves this phase
        }
        // Go on while enough lanes still hold a ray, or while inner nodes are all the block has to do. The test comes AFTER
        // the visit: a claimed ray always advances before it is handed back (two warps that claimed and returned the same
        // rays without visiting them — each failing its first claim while the other held them — ran in lockstep for ever).
        const int held = __popc(__ballot_sync(FULL, row >= 0 && !done));
        if (held == 0 || (held < QB_BP_MIN_LANES && !alone)) break;
      }
      if (row >= 0) {  // whatever this lane still holds goes back to the pool, in the phase it is in now
        ST(F_CUR) = cur;
        ST(F_SP) = (uint32_t) sp;
        ST(F_FLAGS) = flags;
        bp_release(P, bp_phase_of(cur), lane, row);
      }
    } else if (w_leaf >= w_next && w_leaf >= w_fetch) {
      // ---- LEAF: the leaves' triangles, pooled over the warp (see traverse.cuh) --------------------------------------
#ifdef QB_BP_DEBUG
      ++it_leaf;
#endif
      const int row = bp_claim(P, BP_LEAF, lane, aL, start);
      const bool have = row >= 0;
      uint32_t cur = 0u;
      __threadfence_block();  // every lane: the pair loop below reads the owners' slots
      if (have) cur = ST(F_CUR);
      const uint32_t lbase = cur & 0x0FFFFFFFu;
      uint32_t count = have ? ((cur >> 28) & 7u) + 1u : 0u;
      uint32_t incl = count;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += v;
      }
      const bool later = incl > QB_COOP_SLOTS;  // leaves of up to 8 triangles may not all fit one round: back to the pool
      if (later) count = 0u;
      const uint32_t excl = incl - count;
      const uint32_t total = min(__reduce_max_sync(FULL, count ? incl : 0u), (uint32_t) QB_COOP_SLOTS);
      uint32_t *scr = P.coop;
      for (uint32_t i = 0; i < count; ++i) scr[excl + i] = ((uint32_t) lane << 3) | i;
      __syncwarp();
      for (uint32_t q0 = 0; q0 < total; q0 += 32u) {
        const uint32_t q = q0 + (uint32_t) lane;
        const bool mine = q < total;
        const uint32_t code = mine ? scr[q] : 0u;
        const int o = (int) (code >> 3);
        const uint32_t tb = __shfl_sync(FULL, lbase, o);
        const int orow = __shfl_sync(FULL, row, o);
        if (mine) {
          // the owner's ray straight from its slot (column o: lanes on the same owner read the same words)
          const uint32_t *sp0 = &P.state[(F_RPX * P.rows + orow) * 32 + o];
          const int fs = P.rows * 32;
          Ray3 rr;
          rr.p = v3(__uint_as_float(sp0[0]), __uint_as_float(sp0[fs]), __uint_as_float(sp0[2 * fs]));
          rr.d = v3(__uint_as_float(sp0[3 * fs]), __uint_as_float(sp0[4 * fs]), __uint_as_float(sp0[5 * fs]));
          const DTri *tp = &S.tris[tb + (code & 7u)];
          const float4 ta = __ldg(&tp->a), tb4 = __ldg(&tp->b), tc = __ldg(&tp->c), td = __ldg(&tp->d);
          QB_CNT(tris);
          float t;
          const bool ok = tri_candidate(rr, xyz(ta), xyz(tb4), xyz(tc), v3(ta.w, tb4.w, tc.w), td.x, __float_as_int(td.y), t);
          // accepted candidates have t > bias > 0; bit 0 of a rejected entry's word is never looked at
          scr[q] = ok ? __float_as_uint(t) : 0x80000000u;
        }
      }
      __syncwarp();
      if (have && later) bp_release(P, BP_LEAF, lane, row);
      if (have && !later) {
        float z = STF(F_Z);
        int sp = (int) ST(F_SP);
        int obj_slot = -1, uv_slot = -1;
        float uv_t = 0.f;
        bool occluded = false;
        for (uint32_t i = 0; i < count; ++i) {
          const uint32_t w = scr[excl + i];
          const float t = __uint_as_float(w);
          if (!(w & 0x80000000u) && z > t) {
            if (ANY) { occluded = true; break; }
            z = t;
            obj_slot = (int) (lbase + i);
            if (__ldg(&S.tri_face[lbase + i]) & 0x80000000u) { uv_slot = obj_slot; uv_t = t; }
          }
        }
        if (ANY && occluded) {
          io.store_any(ST(F_IDX), true);
          atomicSub(&P.ctrl[1], 1u);
          bp_release(P, BP_FREE, lane, row);
        } else {
          const uint32_t nxt = sp > 0 ? STK(sp - 1) : QB_REF_EMPTY;
          sp -= sp > 0 ? 1 : 0;
          ST(F_CUR) = nxt;
          ST(F_SP) = (uint32_t) sp;
          if (obj_slot >= 0) {
            ST(F_Z) = __float_as_uint(z);
            ST(F_OBJSLOT) = (uint32_t) obj_slot;
            if (uv_slot >= 0) { ST(F_UVSLOT) = (uint32_t) uv_slot; ST(F_UVT) = __float_as_uint(uv_t); }
          }
          bp_release(P, bp_phase_of(nxt), lane, row);
        }
      }
      __syncwarp();  // the scratch is rewritten by this warp's next LEAF phase
    } else if (w_next >= w_fetch) {
      // ---- NEXT: the object just traversed reports its result, then the next objects of the list are set up -------------
#ifdef QB_BP_DEBUG
      ++it_next;
#endif
      const int row = bp_claim(P, BP_NEXT, lane, aN, start);
      const bool have = row >= 0;
      bool active = have;
      uint32_t idx = 0u, cur = QB_REF_EMPTY;
      Ray3 c0, r;
      c0.p = c0.d = r.p = r.d = v3(0.f);
      V3 drcp = v3(0.f);
      float mesh_exit = 0.f, z_entry = 0.f, z_sane = 0.f, uv_t = 0.f;
      int k = 0, sp = 0, cull_state = 0, obj_slot = -1, uv_slot = -1;
      bool gx = false, gy = false, gz = false;
      HitRec h;
      h.z = 0.f; h.k = -1; h.slot = -1; h.tex_k = -1; h.tex_slot = -1; h.tex_t = 0.f; h.last_tri = -1;
      if (have) {
        __threadfence_block();
        idx = ST(F_IDX);
        c0.p = v3(STF(F_C0PX), STF(F_C0PY), STF(F_C0PZ));
        c0.d = v3(STF(F_C0DX), STF(F_C0DY), STF(F_C0DZ));
        h.z = STF(F_Z);
        k = (int) ST(F_K);
        const uint32_t flags = ST(F_FLAGS);
        cull_state = (int) ((flags >> 3) & 3u);
        z_entry = STF(F_ZENTRY);
        z_sane = STF(F_ZSANE);
        obj_slot = (int) ST(F_OBJSLOT);
        uv_slot = (int) ST(F_UVSLOT);
        uv_t = STF(F_UVT);
        if (!ANY) {
          h.k = (int) ST(F_HK); h.slot = (int) ST(F_HSLOT); h.tex_k = (int) ST(F_HTEXK); h.tex_slot = (int) ST(F_HTEXSLOT);
          h.tex_t = STF(F_HTEXT); h.last_tri = (int) ST(F_HLASTTRI);
        }
      }
      bool entered = false;  // a mesh was entered: r, drcp and the per-mesh fields are new
#pragma unroll 1
      for (int rep = 0; rep < QB_NEXT_REPS; ++rep) {
        const bool go_next = active && cur == QB_REF_EMPTY;
        if (rep > 0 && 2 * __popc(__ballot_sync(FULL, go_next)) < n_next) break;
        bool redo = false;
        if (go_next && k > 0 && cull_state == 1) {
          // the cull dropped boxes and the mesh produced no hit inside its own box: the pair is checked exactly, by
          // k_redo_detect / k_redo_exact for the closest-hit rays of the wavefront, else by walking the mesh again here
          cull_state = 0;
          if (!(h.z <= z_sane) && (ANY || !io.defer(idx, k - 1, z_entry))) {
            redo = true;
            cull_state = 2;
            h.z = z_entry;
            obj_slot = -1; uv_slot = -1; sp = 0;
            cur = __float_as_uint(obj_rec(S, T, k - 1).lo.w);
            // r, drcp, mesh_exit and the axis flags of this mesh are still in the slot
          }
        }
        if (go_next && !redo) {
          if (k > 0 && obj_slot >= 0) {
            h.k = k - 1; h.slot = obj_slot; h.last_tri = obj_slot;
            if (uv_slot >= 0) { h.tex_k = k - 1; h.tex_slot = uv_slot; h.tex_t = uv_t; }
            obj_slot = -1;
          }
          if (k == S.num_obj) {
            if (ANY) io.store_any(idx, false); else io.store(idx, h);
            active = false;
          } else {
            const ObjRec o = obj_rec(S, T, k);
            r = chain_ray_tab(S, T, o, k, c0);
            const int type = o.id.x;
            QB_CNT(objects);
            ++k;
            obj_slot = -1; uv_slot = -1;
            if (type == QG_OBJ_MESH) {
              const uint32_t root_ref = __float_as_uint(o.lo.w);
              if (root_ref != QB_NO_FACES && mesh_box_hit_q(o, r, h.z, mesh_exit, drcp)) {
                cur = root_ref;
                sp = 0;
                cull_state = io.exact(idx, k - 1) ? 2 : 0;  // k was advanced above: k - 1 is this object
                z_entry = h.z;
                z_sane = qmax(mesh_exit * 1.001f, o.hi.w);
                gx = qabs(r.d.x) < 1e-7f; gy = qabs(r.d.y) < 1e-7f; gz = qabs(r.d.z) < 1e-7f;
                entered = true;
              }
            } else {
              const float t = type == QG_OBJ_SPHERE ? sphere_t(r, h.z) : plane_t(r, h.z);
              if (t > 0.f) {
                if (ANY) {
                  io.store_any(idx, true);
                  active = false;
                } else {
                  h.z = t; h.k = k - 1; h.slot = -1;
                  h.tex_k = k - 1; h.tex_slot = -1; h.tex_t = t;  // sphere / plane always write uvw
                }
              }
            }
          }
        }
      }
      if (have) {
#ifdef QB_BP_DEBUG
        ST(F_DBG_NEXTS) = ST(F_DBG_NEXTS) + 1u;
#endif
        if (!active) {
          atomicSub(&P.ctrl[1], 1u);
          bp_release(P, BP_FREE, lane, row);
        } else {
          if (entered) {
            ST(F_RPX) = __float_as_uint(r.p.x); ST(F_RPY) = __float_as_uint(r.p.y); ST(F_RPZ) = __float_as_uint(r.p.z);
            ST(F_RDX) = __float_as_uint(r.d.x); ST(F_RDY) = __float_as_uint(r.d.y); ST(F_RDZ) = __float_as_uint(r.d.z);
            ST(F_DCX) = __float_as_uint(drcp.x); ST(F_DCY) = __float_as_uint(drcp.y); ST(F_DCZ) = __float_as_uint(drcp.z);
            ST(F_MEXIT) = __float_as_uint(mesh_exit);
            ST(F_ZENTRY) = __float_as_uint(z_entry);
            ST(F_ZSANE) = __float_as_uint(z_sane);
            ST(F_FLAGS) = (gx ? 1u : 0u) | (gy ? 2u : 0u) | (gz ? 4u : 0u) | ((uint32_t) cull_state << 3);
          } else {
            ST(F_FLAGS) = (ST(F_FLAGS) & 7u) | ((uint32_t) cull_state << 3);
          }
          ST(F_Z) = __float_as_uint(h.z);
          ST(F_CUR) = cur;
          ST(F_SP) = (uint32_t) sp;
          ST(F_K) = (uint32_t) k;
          ST(F_OBJSLOT) = (uint32_t) obj_slot;
          ST(F_UVSLOT) = (uint32_t) uv_slot;
          ST(F_UVT) = __float_as_uint(uv_t);
          if (!ANY) {
            ST(F_HK) = (uint32_t) h.k; ST(F_HSLOT) = (uint32_t) h.slot; ST(F_HTEXK) = (uint32_t) h.tex_k;
            ST(F_HTEXSLOT) = (uint32_t) h.tex_slot; ST(F_HTEXT) = __float_as_uint(h.tex_t); ST(F_HLASTTRI) = (uint32_t) h.last_tri;
          }
          bp_release(P, bp_phase_of(cur), lane, row);
        }
      }
    } else {
      // ---- FETCH: free slots take consecutive rays from the device cursor (one atomic per warp) -----------------------
#ifdef QB_BP_DEBUG
      ++it_fetch;
#endif
      const int row = bp_claim(P, BP_FREE, lane, aF, start);
      const bool have = row >= 0;
      const uint32_t m = __ballot_sync(FULL, have);
      const uint32_t nf = (uint32_t) __popc(m);
      uint32_t base = 0;
      if (lane == 0 && nf) base = atomicAdd(io.cursor(), nf);
      base = __shfl_sync(FULL, base, 0);
      bool got = false;
      if (have) {
        const uint32_t my = base + (uint32_t) __popc(m & lt_mask);
        float tmax = QB_BIGFLOAT;
        Ray3 c;
        const uint32_t ray = my < n ? io.map(my) : my;
        if (my < n && io.load(ray, c, tmax)) {
          const Ray3 c0 = to_node_tab(S, T, 0, c);
          ST(F_IDX) = ray;
          ST(F_C0PX) = __float_as_uint(c0.p.x); ST(F_C0PY) = __float_as_uint(c0.p.y); ST(F_C0PZ) = __float_as_uint(c0.p.z);
          ST(F_C0DX) = __float_as_uint(c0.d.x); ST(F_C0DY) = __float_as_uint(c0.d.y); ST(F_C0DZ) = __float_as_uint(c0.d.z);
          ST(F_Z) = __float_as_uint(tmax);
          ST(F_CUR) = QB_REF_EMPTY;
          ST(F_SP) = 0u;
          ST(F_K) = 0u;
          ST(F_FLAGS) = 0u;
#ifdef QB_BP_DEBUG
          ST(F_DBG_VISITS) = 0u; ST(F_DBG_NEXTS) = 0u;
#endif
          ST(F_ZENTRY) = 0u; ST(F_ZSANE) = 0u; ST(F_MEXIT) = 0u;
          ST(F_OBJSLOT) = 0xFFFFFFFFu; ST(F_UVSLOT) = 0xFFFFFFFFu; ST(F_UVT) = 0u;
          if (!ANY) {
            ST(F_HK) = 0xFFFFFFFFu; ST(F_HSLOT) = 0xFFFFFFFFu; ST(F_HTEXK) = 0xFFFFFFFFu; ST(F_HTEXSLOT) = 0xFFFFFFFFu;
            ST(F_HTEXT) = 0u; ST(F_HLASTTRI) = 0xFFFFFFFFu;
          }
          got = true;
        }
      }
      const uint32_t ng = (uint32_t) __popc(__ballot_sync(FULL, got));
      if (lane == 0) {
        if (ng) atomicAdd(&P.ctrl[1], ng);  // alive BEFORE the source may be declared exhausted
        __threadfence_block();
        if (nf && base + nf >= n) *(volatile uint32_t *) &P.ctrl[0] = 1u;
      }
      __syncwarp();
      if (have) bp_release(P, got ? BP_NEXT : BP_FREE, lane, row);
    }
  }
}

#undef ST
#undef STF
#undef STK

}  // namespace qb
