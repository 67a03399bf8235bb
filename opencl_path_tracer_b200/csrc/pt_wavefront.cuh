// pt_wavefront.cuh — wavefront organisation of the same path (included inside namespace PT_NS).
//
// One PATH per pixel lives in global memory (ray 32 B + state 80-88 B, of which a path step touches 48-56 B: wf_store_hot).  A render is a sequence of
// iterations; each iteration advances every active path by one ray:
//     k_wf_trace   persistent warps; every lane pulls the next queued path as soon as its traversal
//                  ends (warp-aggregated atomic fetch), so no lane waits for a neighbour's long
//                  traversal; inside, box steps and leaf steps are phase-scheduled by warp votes
//     k_wf_shade   one thread per queued path, full warps: hit → material → bounce sampling →
//                  throughput update (prog.cl:321-366), path end → running mean (prog.cl:379) and the
//                  NEXT sample's camera ray (prog.cl:384-389) in the same thread; the new ray's scan over the
//                  LBVH's outlier list (its result travels in the ray record as the traversal's initial best
//                  hit); surviving paths are appended to the next queue with a warp ballot + one atomic per warp
// The frame's paths are dealt to 1..4 independent sub-wavefronts (own queues, own stream) whose kernels fill each
// other's tails (launch_wavefront).  k_wf_trace walks the 4-ary collapse of the LBVH: per node 3 texture fetches (x
// rows by direction sign, links) + 2 256-bit loads (y, z row pairs), per leaf triangle 2 256-bit loads of a 64-byte
// record — the L1 data pipes (one wavefront per lane and scattered access), not DRAM, are what the traversal loads.
// A pixel's samples stay strictly sequential (same RNG order as the megakernel and the reference),
// so the result is bit-identical to the other pipelines in the parity flavour.
// Iterations needed: at most spp * iterations (every active path consumes one ray per iteration).

#define WF_PLANES 6                      // float4 planes of W*H reserved per frame (ptb_abi.cu, pt_pool.cuh size their scratch by it)
struct WfPath {
    Path ps;
    V3 acc;
    uint32_t rng;
    int s, depth;
};
// Path state in global memory, split by HOW OFTEN a path step needs it (round-1 verdict: 96 B read + 96 B written per step).
//   hot   (every step): the throughput factors fL, fB, fS, the RNG state, depth | inside-glass flag — three float4 planes, 48 B.
//         The parity flavour also keeps fR (the fast one folds it into fS, shade_hit): its x in the spare word, y z in a float2 plane.
//   tally (sample end only, ~1 step in 3): running mean / sum of the pixel and the sample index — one float4 plane.
//   color (sample end, emitter hits, flat preview): the sample's radiance so far — one float4 plane.
// Planes 0-2 hot, 3 tally, 4 color, first half of 5: the parity flavour's float2 plane.
PT_DEV void wf_store_hot(const RenderParams &p, int i, const WfPath &w) {
    const size_t n = (size_t)p.W * p.H;
    float4 *st = p.wf_state + i;
    const int di = w.depth | (w.ps.in ? (int)0x80000000 : 0);
    st[0 * n] = make_float4(w.ps.fL.x, w.ps.fL.y, w.ps.fL.z, w.ps.fB.x);
    st[1 * n] = make_float4(w.ps.fB.y, w.ps.fB.z, w.ps.fS.x, w.ps.fS.y);
#if PT_PARITY
    st[2 * n] = make_float4(w.ps.fS.z, __uint_as_float(w.rng), __int_as_float(di), w.ps.fR.x);
    reinterpret_cast<float2 *>(p.wf_state + 5 * n)[i] = make_float2(w.ps.fR.y, w.ps.fR.z);
#else
    st[2 * n] = make_float4(w.ps.fS.z, __uint_as_float(w.rng), __int_as_float(di), 0.0f);
#endif
}
PT_DEV void wf_load_hot(const RenderParams &p, int i, WfPath &w) {
    const size_t n = (size_t)p.W * p.H;
    const float4 *st = p.wf_state + i;
    const float4 a = st[0 * n], b = st[1 * n], c = st[2 * n];
    w.ps.fL = v3(a.x, a.y, a.z); w.ps.fB = v3(a.w, b.x, b.y); w.ps.fS = v3(b.z, b.w, c.x);
    w.rng = __float_as_uint(c.y);
    const int di = __float_as_int(c.z);
    w.depth = di & 0x7fffffff; w.ps.in = di < 0;
#if PT_PARITY
    const float2 r = reinterpret_cast<const float2 *>(p.wf_state + 5 * n)[i];
    w.ps.fR = v3(c.w, r.x, r.y);
#else
    w.ps.fR = v3(1.0f, 1.0f, 1.0f);
#endif
}
PT_DEV void wf_store_tally(const RenderParams &p, int i, const WfPath &w) {
    p.wf_state[3 * (size_t)p.W * p.H + i] = make_float4(w.acc.x, w.acc.y, w.acc.z, __int_as_float(w.s));
}
PT_DEV void wf_load_tally(const RenderParams &p, int i, WfPath &w) {
    const float4 e = p.wf_state[3 * (size_t)p.W * p.H + i];
    w.acc = v3(e.x, e.y, e.z); w.s = __float_as_int(e.w);
}
PT_DEV void wf_store_color(const RenderParams &p, int i, const WfPath &w) {
    p.wf_state[4 * (size_t)p.W * p.H + i] = make_float4(w.ps.color.x, w.ps.color.y, w.ps.color.z, 0.0f);
}
PT_DEV void wf_load_color(const RenderParams &p, int i, WfPath &w) {
    const float4 d = p.wf_state[4 * (size_t)p.W * p.H + i];
    w.ps.color = v3(d.x, d.y, d.z);
}
PT_DEV void wf_store_ray(const RenderParams &p, int i, V3 P, V3 D, const Closest &best) {
    p.wf_ray0[i] = make_float4(P.x, P.y, P.z, D.x);
    p.wf_ray1[i] = make_float4(D.y, D.z, best.t, __int_as_float(best.tri));
}

// The LBVH's outlier list (room walls, emitter quads: lbvh.cu step 2) is scanned by the kernel that CREATES a ray — k_wf_init /
// k_wf_shade, full warps — and its result travels in the ray record as the traversal's initial best hit.  Inside the trace
// kernel the same scan ran whenever a few lanes refilled, at 11 of 32 lanes (profiles/r01_wf_trace_v6.txt: 15 % of its
// instructions).  Fast flavour: the outliers as merged quads in the pair layout of brute_smem_prims (pt_kernels.cuh), staged
// in shared memory by the block; parity flavour and the counting build: the plain triangles.
#define WF_BIG_SMEM (7 * (LBVH_BIG_MAX / 2))
PT_DEV int wf_stage_big(const RenderParams &p, float4 *s_big) {
#if !PT_PARITY
    const SceneDev &sc = p.sc;
    const int n_big_pairs = p.accel != PT_ACCEL_LBVH ? 0 : (sc.n_big_prims + 1) >> 1;
    for (int j = threadIdx.x; j < n_big_pairs; j += blockDim.x) {
        const int ia = 2 * j, ib = 2 * j + 1;
        const bool hb = ib < sc.n_big_prims;
        const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        const float4 A0 = ldg4(sc.bq0 + ia), A1 = ldg4(sc.bq1 + ia), A2 = ldg4(sc.bq2 + ia);
        const float4 B0 = hb ? ldg4(sc.bq0 + ib) : z, B1 = hb ? ldg4(sc.bq1 + ib) : z, B2 = hb ? ldg4(sc.bq2 + ib) : z;
        const float2 sa = __ldg(sc.big_ab + ia), sb = hb ? __ldg(sc.big_ab + ib) : make_float2(1.0f, 0.0f);
        s_big[7 * j + 0] = make_float4(A0.x, B0.x, A0.y, B0.y);
        s_big[7 * j + 1] = make_float4(A0.z, B0.z, -A0.w, -B0.w);
        s_big[7 * j + 2] = make_float4(-A1.x, -B1.x, -A1.y, -B1.y);
        s_big[7 * j + 3] = make_float4(A1.z, B1.z, A1.w, B1.w);
        s_big[7 * j + 4] = make_float4(A2.x, B2.x, A2.y, B2.y);
        s_big[7 * j + 5] = make_float4(A2.z, B2.z, A2.w, B2.w);
        s_big[7 * j + 6] = make_float4(-sa.x, -sb.x, -sa.y, -sb.y);
    }
    __syncthreads();
    return n_big_pairs;
#else
    return 0;
#endif
}
PT_DEV void wf_big_scan(const RenderParams &p, const float4 *s_big, int n_big_pairs, V3 P, V3 D, Closest &best, Work &wk) {
    best.t = PT_NO_HIT; best.tri = 0x7fffffff;
    if (p.accel != PT_ACCEL_LBVH) return;
#if !PT_PARITY
    if (n_big_pairs > 0) {                            // packed scan over the merged outliers, then primitive -> upload index
        if (p.count_work) brute_smem_prims<2, true>(s_big, n_big_pairs, P, D, best, &wk, p.sc.n_big_prims);
        else brute_smem_prims<2>(s_big, n_big_pairs, P, D, best);
        if (best.t < PT_NO_HIT) best.tri = __ldg(p.sc.big_rep + best.tri);
        return;
    }
#endif
    if (p.count_work) lbvh_big_scan<true>(p.sc, P, D, best, wk);
    else lbvh_big_scan<false>(p.sc, P, D, best, wk);
}

// append `i` to the output queue for the lanes with keep == true: one atomic per warp
PT_DEV void wf_push(int *queue, int *count, bool keep, int i) {
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (m == 0u) return;
    const int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == (__ffs(m) - 1)) base = atomicAdd(count, __popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    if (keep) queue[base + __popc(m & ((1u << lane) - 1u))] = i;
}

// The same for a whole block, grouped by direction octant: the block's survivors are appended as one contiguous run in
// which rays of one octant sit together (a counting sort over 8 buckets in shared memory), so the 32 rays a trace warp
// fetches mostly descend the tree in the same child order.  Origins stay local (one block = 256 neighbouring paths).
PT_DEV void wf_push_by_octant(int *queue, int *count, bool keep, int i, V3 D) {
    __shared__ int s_cnt[8], s_off[8];
    if (threadIdx.x < 8) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int oct = (D.x < 0.0f ? 1 : 0) | (D.y < 0.0f ? 2 : 0) | (D.z < 0.0f ? 4 : 0);
    const unsigned peers = __match_any_sync(0xffffffffu, keep ? oct : 8);
    int rank = 0;
    if (keep) {
        const int leader = __ffs(peers) - 1;
        if (lane == leader) rank = atomicAdd(&s_cnt[oct], __popc(peers));
        rank = __shfl_sync(peers, rank, leader) + __popc(peers & ((1u << lane) - 1u));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int tot = 0;
        for (int o = 0; o < 8; ++o) { s_off[o] = tot; tot += s_cnt[o]; }
        const int base = tot ? atomicAdd(count, tot) : 0;
        for (int o = 0; o < 8; ++o) s_off[o] += base;
    }
    __syncthreads();
    if (keep) queue[s_off[oct] + rank] = i;
}

// starts the next sample(s) of pixel i until a path is in flight or all samples are done; returns alive
PT_DEV bool wf_next_sample(const RenderParams &p, int i, WfPath &w, V3 &rayP, V3 &rayD, int s_end, Work &wk) {
    const int camX = (int)p.cam.XM;
    while (w.s < s_end) {
        float r1 = pt_rand(w.rng), r2 = pt_rand(w.rng);
        camera_ray(p, i % max(camX, 1), i / max(camX, 1), r1, r2, rayP, rayD);
        path_reset(w.ps);
        w.depth = 0;
        if (p.iterations > 0) return true;
        if (p.sum_mode) w.acc = w.acc + w.ps.color; else w.acc = pt_mean_step(w.acc, w.ps.color, w.s);
        ++w.s;
        wk.samples++;
    }
    return false;
}
PT_DEV void wf_finish_pixel(const RenderParams &p, int i, const WfPath &w) {
    p.rnds[i] = (int)w.rng;
    p.colors[i] = make_float4(w.acc.x, w.acc.y, w.acc.z, 0.0f);
    if (p.tex) p.tex[i] = reinhard_tone_map(w.acc);
}
PT_DEV void wf_flush_counters(const RenderParams &p, const Work &wk, bool count) {
    if (!p.counters) return;
    const int lane = threadIdx.x & 31;
    unsigned v[7] = {wk.rays, wk.box, wk.tri, wk.tri1, wk.tri2, wk.tri3, wk.samples};
    const int idx[7] = {PT_CNT_RAYS, PT_CNT_BOX, PT_CNT_TRI, PT_CNT_TRI1, PT_CNT_TRI2, PT_CNT_TRI3, PT_CNT_PIXEL_SAMPLES};
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        if (!count && k >= 1 && k <= 5) continue;
        unsigned long long tot = v[k];
        for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
        if (lane == 0 && tot) atomicAdd(p.counters + idx[k], tot);
    }
}

// ---------------------------------------------------------------- init: first camera ray of every selected pixel
// `subs` interleaved sub-wavefronts (launch_wavefront): block b feeds sub b % subs, whose first queue starts at
// wf_queue[0] + sub * sub_cap and whose control words are wf_ctrl + 4 * sub
__global__ void __launch_bounds__(256) k_wf_init(const RenderParams p, const int subs, const int sub_cap) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int sub = blockIdx.x % subs;
    const int n = p.W * p.H;
    Work wk = {0, 0, 0, 0, 0, 0, 0};
    __shared__ float4 s_big[WF_BIG_SMEM];
    const int n_big_pairs = wf_stage_big(p, s_big);
    bool keep = false;
    if (i < n && (p.shard_n <= 1 || ((i / p.shard_group) % p.shard_n) == p.shard_rank)) {
        WfPath w;
        w.rng = (uint32_t)p.rnds[i];
        w.acc = (p.sample0 != 0 || p.sum_mode) ? v3(p.colors[i]) : v3(0.0f, 0.0f, 0.0f);
        w.s = p.sample0;
        w.depth = 0;
        path_reset(w.ps);
        V3 P = v3(0.0f, 0.0f, 0.0f), D = v3(0.0f, 0.0f, 1.0f);
        keep = wf_next_sample(p, i, w, P, D, p.sample0 + p.spp, wk);
        if (keep) {
            Closest best;
            wf_big_scan(p, s_big, n_big_pairs, P, D, best, wk);
            wf_store_ray(p, i, P, D, best);
            wf_store_hot(p, i, w);
            wf_store_tally(p, i, w);
            wf_store_color(p, i, w);
        } else wf_finish_pixel(p, i, w);
    }
    wf_push(p.wf_queue[0] + (size_t)sub * sub_cap, p.wf_ctrl + 4 * sub, keep, i);
    wf_flush_counters(p, wk, p.count_work != 0);
}

// ---------------------------------------------------------------- shade: one thread per queued path
// q_in/n_in: this iteration's queue; survivors go to q_out/n_out (zeroed by the host before the launch)
#ifndef WF_SHADE_MINB
#define WF_SHADE_MINB 4            // 64 registers: 32 resident warps per SM against the four dependent loads of a path step
#endif
__global__ void __launch_bounds__(256, WF_SHADE_MINB) k_wf_shade(const RenderParams p, const int *q_in, const int *n_in, int *q_out, int *n_out) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int count = *n_in;
    Work wk = {0, 0, 0, 0, 0, 0, 0};
    __shared__ float4 s_big[WF_BIG_SMEM];
    const int n_big_pairs = wf_stage_big(p, s_big);
    bool keep = false;
    int i = 0;
    V3 rayP = v3(0.0f, 0.0f, 0.0f), rayD = v3(0.0f, 0.0f, 1.0f);
    if (j < count) {
        i = q_in[j];
        const SceneDev &sc = p.sc;
        const float4 r0 = p.wf_ray0[i], r1 = p.wf_ray1[i];
        rayP = v3(r0.x, r0.y, r0.z); rayD = v3(r0.w, r1.x, r1.y);
        const float t = r1.z;
        const int tri = __float_as_int(r1.w);
        WfPath w;
        wf_load_hot(p, i, w);
        const V3 eye = v3(p.cam.eye.x, p.cam.eye.y, p.cam.eye.z);
        const bool hit = t < PT_NO_HIT;
        // the sample ends with this step on a miss (prog.cl:367-376) or at the last bounce: known before the material is,
        // so the planes only a finished sample needs are requested here, beside the triangle's
        const bool ends = !hit || w.depth + 1 >= p.iterations;
        if (ends) { wf_load_tally(p, i, w); wf_load_color(p, i, w); }
        wk.rays++;
        if (hit) {
            V3 hitP = rayP + rayD * t;                               // prog.cl:107
            V3 hitN = tri_normal(ldg4(sc.q0 + tri), ldg4(sc.q1 + tri), ldg4(sc.q2 + tri));
            Mat m = load_material(sc.mats, (int)__ldg(sc.mati + tri));
            const bool lit = m.type == 3 || p.iterations == 1;       // the only steps that touch the sample's radiance (shade_hit)
            if (lit && !ends) wf_load_color(p, i, w);
            shade_hit(m, hitP, hitN, rayP, rayD, w.rng, w.ps, eye, p.iterations);
            if (lit && !ends) wf_store_color(p, i, w);
            ++w.depth;
        }
        bool alive = !ends;
        if (ends) {                                                  // sample finished, prog.cl:379
            if (p.sum_mode) w.acc = w.acc + w.ps.color; else w.acc = pt_mean_step(w.acc, w.ps.color, w.s);
            ++w.s;
            wk.samples++;
            alive = wf_next_sample(p, i, w, rayP, rayD, p.sample0 + p.spp, wk);
            if (alive) { wf_store_tally(p, i, w); wf_store_color(p, i, w); }
        }
        if (alive) {
            Closest best;
            wf_big_scan(p, s_big, n_big_pairs, rayP, rayD, best, wk);
            wf_store_ray(p, i, rayP, rayD, best);
            wf_store_hot(p, i, w);
        } else wf_finish_pixel(p, i, w);
        keep = alive;
    }
    if (p.wf_sort) wf_push_by_octant(q_out, n_out, keep, i, rayD);
    else wf_push(q_out, n_out, keep, i);
    wf_flush_counters(p, wk, p.count_work != 0);
}

// ---------------------------------------------------------------- trace: persistent warps over the queue
enum { WF_FETCH = 0, WF_NODE = 1, WF_LEAF = 2, WF_EXIT = 3, WF_POP = 4 };
#ifndef WF_FETCH_MIN
#define WF_FETCH_MIN 8            // refill when this many lanes are idle (or nothing else is runnable)
#endif
#ifndef WF_SMEM_STACK
#define WF_SMEM_STACK 16
#endif

// Lane states: FETCH (needs a ray), NODE (link is an inner node), LEAF (link is a leaf), POP (needs the next stack entry).
// Every warp step votes one phase.  Inside the NODE phase the code is straight-line: the far child is pushed by a
// predicated store, and ALL lanes that need a stack entry (no child hit, or a leaf just finished) share one pop block
// that tries two entries per step and drops the culled ones — the first version branched three ways per lane here and
// spent 47 % of its instructions at fewer than 12 active lanes (profiles/r01_wf_trace_v2.txt).
// WF_POSTPONE: a lane that reaches a leaf while it holds no postponed one parks the leaf link in `pend` and keeps
// traversing (its next stack entry) instead of idling until enough lanes want a leaf step; the leaf step then tests the
// postponed and the current leaf of EVERY lane that has one.  The extra node visits are speculative (they run in slots
// the lane would have idled through, and anything they push is culled at pop once the leaf has shrunk best.t); the
// result cannot change: every leaf whose box passed is still tested before the ray finishes.
// MEASURED (scripts/wf_ab.sh): 3.44 vs 3.48 Grays/s (104 k), 3.12 vs 3.15 (1 M) — box tests per ray +4 %, no time won:
// the lanes the leaf step gains are lost again to the speculative visits.  Off by default, kept as the experiment.
#ifndef WF_POSTPONE
#define WF_POSTPONE 0
#endif
#ifndef WF_TRI_RECORDS
#define WF_TRI_RECORDS 1
#endif
#ifndef WF_SPLIT_FETCH
#define WF_SPLIT_FETCH 3
#endif
#ifndef WF_MINB
#define WF_MINB 8                 // resident blocks per SM the register allocation is held to
#endif
// WIDE: 0 = binary nodes, 1 = 4-ary float nodes (128 bytes), 2 = 4-ary quantised nodes (64 bytes, lbvh.cu k_emit4q)
template <bool COUNT, int WIDE>
__global__ void __launch_bounds__(PT_BLOCK, WF_MINB) k_wf_trace(const RenderParams p, const int *q_in, const int *n_in, int *cursor, const int3 tune) {
    const SceneDev &sc = p.sc;
    const int lane = threadIdx.x & 31;
    const int count = *n_in;
    Work wk = {0, 0, 0, 0, 0, 0, 0};
    __shared__ int2 s_stack[WF_SMEM_STACK * PT_BLOCK];
    int2 *my_stack = s_stack + threadIdx.x;
    int2 deep_stack[PT_LBVH_STACK - WF_SMEM_STACK];
    int st = WF_FETCH, link = 0, sp = 0, path = 0;
    int pend = 0;                                        // postponed leaf link (< 0), 0 = none
    bool exhausted = count == 0;
    V3 P = v3(0.0f, 0.0f, 0.0f), D = v3(0.0f, 0.0f, 1.0f);
    Closest best; best.t = PT_NO_HIT; best.tri = 0x7fffffff;
    float ix = 0.0f, iy = 0.0f, iz = 0.0f, ox = 0.0f, oy = 0.0f, oz = 0.0f;
    int nx = 0, ny = 0, nz = 0;                          // row of the NEAR plane per axis: 0 = lo (direction >= 0), 1 = hi
#ifdef WF_DEBUG_STATES
    unsigned dbgN = 0, dbgL = 0, dbgF = 0, dbgP = 0, dbgI = 0;
#endif
#define WF_PUSH(v, d) do { const int2 e_ = make_int2((v), __float_as_int(d)); if (sp < WF_SMEM_STACK) my_stack[sp * PT_BLOCK] = e_; \
                           else if (sp < PT_LBVH_STACK) deep_stack[sp - WF_SMEM_STACK] = e_; ++sp; } while (0)
#define WF_FINISH() do { float4 *o_ = p.wf_ray1 + path; o_->z = best.t; o_->w = __int_as_float(best.tri); st = WF_FETCH; } while (0)
    for (;;) {
        if (exhausted && st == WF_FETCH) st = WF_EXIT;
        const unsigned bF = __ballot_sync(0xffffffffu, st == WF_FETCH);
        const unsigned bN = __ballot_sync(0xffffffffu, st == WF_NODE || st == WF_POP);
        const unsigned bL = __ballot_sync(0xffffffffu, st == WF_LEAF);
        if ((bF | bN | bL) == 0u) break;
        const int cF = __popc(bF), cN = __popc(bN), cL = __popc(bL);
#ifdef WF_DEBUG_STATES   // throw-away instrumentation: lane-state census per vote, reported through the work counters
        { const int cP = __popc(__ballot_sync(0xffffffffu, st == WF_POP)); dbgN += cN; dbgL += cL; dbgF += cF; dbgP += cP; dbgI += 1; }
#endif
        // tune.x: idle lanes that trigger a refill, tune.y: lanes at a leaf that trigger a leaf step, tune.z: node steps per vote
        if (cF > 0 && (cF >= tune.x || (bN | bL) == 0u)) {
            // ------------------------------------------------------------ FETCH: one atomic for the whole warp
            int base = 0;
            if (lane == (__ffs(bF) - 1)) base = atomicAdd(cursor, cF);
            base = __shfl_sync(0xffffffffu, base, __ffs(bF) - 1);
            if (base + cF >= count) exhausted = true;
            if (st == WF_FETCH) {
                const int j = base + __popc(bF & ((1u << lane) - 1u));
                if (j < count) {
                    path = q_in[j];
                    const float4 r0 = p.wf_ray0[path], r1 = p.wf_ray1[path];
                    P = v3(r0.x, r0.y, r0.z); D = v3(r0.w, r1.x, r1.y);
                    ix = safe_rcp(D.x); iy = safe_rcp(D.y); iz = safe_rcp(D.z);
                    if (WIDE == 2) {                                   // bounded, so that 128 * step * (1 / d) of the quantised decode cannot overflow
                        ix = fminf(fmaxf(ix, -1e25f), 1e25f); iy = fminf(fmaxf(iy, -1e25f), 1e25f); iz = fminf(fmaxf(iz, -1e25f), 1e25f);
                    }
                    ox = P.x * ix; oy = P.y * iy; oz = P.z * iz;
                    nx = ix < 0.0f ? 1 : 0; ny = iy < 0.0f ? 1 : 0; nz = iz < 0.0f ? 1 : 0;
                    best.t = r1.z; best.tri = __float_as_int(r1.w);        // the outlier scan ran where the ray was created (wf_big_scan)
                    sp = 0;
                    pend = 0;
                    link = sc.bvh_root;
                    st = link >= 0 ? WF_NODE : WF_LEAF;
                } else {
                    st = WF_EXIT;
                }
            }
        } else if (cN > 0 && cL < tune.y) {
            // ------------------------------------------------------------ NODE steps
#pragma unroll 1
            for (int k = 0; k < tune.z; ++k) {
                // pop FIRST: a lane that gets a node link here runs its box tests in this same step
                if (st == WF_POP) {
#pragma unroll
                    for (int attempt = 0; attempt < 2; ++attempt) {
                        if (st == WF_POP) {
                            if (sp == 0) {
                                if (WF_POSTPONE && pend != 0) { link = 0; st = WF_LEAF; }     // only the postponed leaf is left
                                else WF_FINISH();
                            } else {
                                --sp;
                                const int2 e = sp < WF_SMEM_STACK ? my_stack[sp * PT_BLOCK] : deep_stack[sp - WF_SMEM_STACK];
                                if (__int_as_float(e.y) <= best.t) {
                                    if (WF_POSTPONE && e.x < 0 && pend == 0) pend = e.x;      // park it, keep popping
                                    else { link = e.x; st = e.x >= 0 ? WF_NODE : WF_LEAF; }
                                }
                            }
                        }
                    }
                }
                if (WIDE == 2 && st == WF_NODE) {
                    // quantised wide node: 64 bytes = four texture fetches (or two 256-bit loads) instead of the float node's five fetches of
                    // 128 bytes; a bound byte q becomes the float 256 + q / 128 with one PRMT, the plane distance one FFMA (see k_pool, pt_pool.cuh)
                    unsigned w0, w1, w2, w3, w4, w5, w6, w7, w8, w9, k0u, k1u, k2u, k3u;
                    if (sc.bvh4q_tex) {
                        const uint4 a0 = tex1Dfetch<uint4>(sc.bvh4q_tex, 4 * link), a1 = tex1Dfetch<uint4>(sc.bvh4q_tex, 4 * link + 1);
                        const uint4 a2 = tex1Dfetch<uint4>(sc.bvh4q_tex, 4 * link + 2), a3 = tex1Dfetch<uint4>(sc.bvh4q_tex, 4 * link + 3);
                        w0 = a0.x; w1 = a0.y; w2 = a0.z; w3 = a0.w; w4 = a1.x; w5 = a1.y; w6 = a1.z; w7 = a1.w;
                        w8 = a2.x; w9 = a2.y; k0u = a3.x; k1u = a3.y; k2u = a3.z; k3u = a3.w;
                    } else {
                        const uint4 *nq = reinterpret_cast<const uint4 *>(sc.bvh4q) + 4 * (size_t)link;
                        unsigned w10, w11;
                        asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                     : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3), "=r"(w4), "=r"(w5), "=r"(w6), "=r"(w7) : "l"(nq));
                        asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                     : "=r"(w8), "=r"(w9), "=r"(w10), "=r"(w11), "=r"(k0u), "=r"(k1u), "=r"(k2u), "=r"(k3u) : "l"(nq + 2));
                        (void)w10; (void)w11;
                    }
                    if (COUNT) wk.box += 4;
                    const float Ax = __uint_as_float((w3 & 0xffu) << 23) * ix, Ay = __uint_as_float(((w3 >> 8) & 0xffu) << 23) * iy,
                                Az = __uint_as_float(((w3 >> 16) & 0xffu) << 23) * iz;
                    const float Cx = fmaf(-256.0f, Ax, fmaf(__uint_as_float(w0), ix, -ox)), Cy = fmaf(-256.0f, Ay, fmaf(__uint_as_float(w1), iy, -oy)),
                                Cz = fmaf(-256.0f, Az, fmaf(__uint_as_float(w2), iz, -oz));
                    const unsigned nxw = nx ? w5 : w4, fxw = nx ? w4 : w5, nyw = ny ? w7 : w6, fyw = ny ? w6 : w7, nzw = nz ? w9 : w8, fzw = nz ? w8 : w9;
                    float d0, d1, d2, d3;
#define WF_QF(word, sel) __uint_as_float(__byte_perm((word), 0x43800000u, (sel)))
#define WF_QSLAB(dk, sel) do { float lo_ = fmaxf(fmaxf(fmaf(WF_QF(nxw, sel), Ax, Cx), fmaf(WF_QF(nyw, sel), Ay, Cy)), fmaf(WF_QF(nzw, sel), Az, Cz));   \
                               float hi_ = fminf(fminf(fmaf(WF_QF(fxw, sel), Ax, Cx), fmaf(WF_QF(fyw, sel), Ay, Cy)), fmaf(WF_QF(fzw, sel), Az, Cz));   \
                               lo_ = fmaxf(lo_ * 0.999996f, 0.0f); hi_ = fminf(hi_ * 1.000004f, best.t);                                                \
                               dk = lo_ <= hi_ ? lo_ : PT_NO_HIT; } while (0)
                    WF_QSLAB(d0, 0x7604); WF_QSLAB(d1, 0x7614); WF_QSLAB(d2, 0x7624); WF_QSLAB(d3, 0x7634);
#undef WF_QSLAB
#undef WF_QF
                    int k0 = (int)k0u, k1 = (int)k1u, k2 = (int)k2u, k3 = (int)k3u;
#define WF_CE(da, ka, db, kb) do { const bool sw_ = db < da; const float td_ = sw_ ? db : da; const int tk_ = sw_ ? kb : ka; \
                                   db = sw_ ? da : db; kb = sw_ ? ka : kb; da = td_; ka = tk_; } while (0)
                    WF_CE(d0, k0, d1, k1); WF_CE(d2, k2, d3, k3); WF_CE(d0, k0, d2, k2); WF_CE(d1, k1, d3, k3); WF_CE(d1, k1, d2, k2);
#undef WF_CE
                    if (d3 < PT_NO_HIT) WF_PUSH(k3, d3);
                    if (d2 < PT_NO_HIT) WF_PUSH(k2, d2);
                    if (d1 < PT_NO_HIT) WF_PUSH(k1, d1);
                    if (WF_POSTPONE && d0 < PT_NO_HIT && k0 < 0 && pend == 0) { pend = k0; st = WF_POP; }
                    else { link = k0; st = d0 < PT_NO_HIT ? (k0 >= 0 ? WF_NODE : WF_LEAF) : WF_POP; }
                }
                if (WIDE == 1 && st == WF_NODE) {
                    // 4-ary node (lbvh.cu): four slab tests, hits sorted by entry distance, nearest visited, rest pushed far-to-near.
                    // The near / far plane of every axis is picked by the SIGN of the ray direction when the row is fetched
                    // (rows 0/1 = lo.x/hi.x of the four children, 2/3 = y, 4/5 = z), not by a min/max per child and axis:
                    // fma(X, ix, -ox) is monotonic in X, so the picked values ARE min(a, b) / max(a, b) of the old test.
                    // Hit <=> max(entry * (1 - 4e-6), 0) <= min(exit * (1 + 4e-6), best.t); the clamped entry is the sort / cull key.
                    const float4 *nd = sc.bvh4 + 8 * (size_t)link;
                    float4 Xn, Xf, Yn, Yf, Zn, Zf, LK;
                    if (sc.bvh4_tex) {
                        // The node's seven rows are split over BOTH data pipes of the L1: x and y rows through the texture pipe,
                        // the z row pair (one 256-bit load) and the links through the LSU pipe.  All seven through the
                        // texture pipe kept it 76 % busy (l1tex__data_pipe_tex_wavefronts, profiles/r01_wf_trace_v7.txt:
                        // one wavefront per lane and fetch) while the LSU pipe sat at 27 %.
                        const int t_ = 8 * link;
                        // WF_SPLIT_FETCH bits: 1 = z rows, 2 = y rows, 4 = x rows as one 256-bit LSU load; 8 = links through the LSU
                        if (WF_SPLIT_FETCH & 4) { float4 A0, A1; ldg8(nd, A0, A1); Xn = nx ? A1 : A0; Xf = nx ? A0 : A1; }
                        else { Xn = tex1Dfetch<float4>(sc.bvh4_tex, t_ + nx); Xf = tex1Dfetch<float4>(sc.bvh4_tex, t_ + (nx ^ 1)); }
                        if (WF_SPLIT_FETCH & 2) { float4 A0, A1; ldg8(nd + 2, A0, A1); Yn = ny ? A1 : A0; Yf = ny ? A0 : A1; }
                        else { Yn = tex1Dfetch<float4>(sc.bvh4_tex, t_ + 2 + ny); Yf = tex1Dfetch<float4>(sc.bvh4_tex, t_ + 2 + (ny ^ 1)); }
                        if (WF_SPLIT_FETCH & 1) { float4 A0, A1; ldg8(nd + 4, A0, A1); Zn = nz ? A1 : A0; Zf = nz ? A0 : A1; }
                        else { Zn = tex1Dfetch<float4>(sc.bvh4_tex, t_ + 4 + nz); Zf = tex1Dfetch<float4>(sc.bvh4_tex, t_ + 4 + (nz ^ 1)); }
                        if (WF_SPLIT_FETCH & 8) LK = ldg4(nd + 6);
                        else LK = tex1Dfetch<float4>(sc.bvh4_tex, t_ + 6);
                    } else {
                        Xn = ldg4(nd + nx); Xf = ldg4(nd + (nx ^ 1)); Yn = ldg4(nd + 2 + ny); Yf = ldg4(nd + 2 + (ny ^ 1));
                        Zn = ldg4(nd + 4 + nz); Zf = ldg4(nd + 4 + (nz ^ 1)); LK = ldg4(nd + 6);
                    }
                    if (COUNT) wk.box += 4;
                    float d0, d1, d2, d3;
#define WF_SLAB(dk, c) do { float lo_ = fmaxf(fmaxf(fmaf(Xn.c, ix, -ox), fmaf(Yn.c, iy, -oy)), fmaf(Zn.c, iz, -oz));                 \
                            float hi_ = fminf(fminf(fmaf(Xf.c, ix, -ox), fmaf(Yf.c, iy, -oy)), fmaf(Zf.c, iz, -oz));                 \
                            lo_ = fmaxf(lo_ * 0.999996f, 0.0f); hi_ = fminf(hi_ * 1.000004f, best.t);                                \
                            dk = lo_ <= hi_ ? lo_ : PT_NO_HIT; } while (0)
                    WF_SLAB(d0, x); WF_SLAB(d1, y); WF_SLAB(d2, z); WF_SLAB(d3, w);
#undef WF_SLAB
                    int k0 = __float_as_int(LK.x), k1 = __float_as_int(LK.y), k2 = __float_as_int(LK.z), k3 = __float_as_int(LK.w);
#define WF_CE(da, ka, db, kb) do { const bool sw_ = db < da; const float td_ = sw_ ? db : da; const int tk_ = sw_ ? kb : ka; \
                                   db = sw_ ? da : db; kb = sw_ ? ka : kb; da = td_; ka = tk_; } while (0)
                    WF_CE(d0, k0, d1, k1); WF_CE(d2, k2, d3, k3); WF_CE(d0, k0, d2, k2); WF_CE(d1, k1, d3, k3); WF_CE(d1, k1, d2, k2);
#undef WF_CE
                    if (d3 < PT_NO_HIT) WF_PUSH(k3, d3);
                    if (d2 < PT_NO_HIT) WF_PUSH(k2, d2);
                    if (d1 < PT_NO_HIT) WF_PUSH(k1, d1);
                    if (WF_POSTPONE && d0 < PT_NO_HIT && k0 < 0 && pend == 0) { pend = k0; st = WF_POP; }
                    else { link = k0; st = d0 < PT_NO_HIT ? (k0 >= 0 ? WF_NODE : WF_LEAF) : WF_POP; }
                }
                if (WIDE == 0 && st == WF_NODE) {
                    const float4 *nd = sc.bvh + 4 * (size_t)link;
                    float4 n0, n1, n2, n3;
                    if (sc.bvh_tex) {
                        const int t_ = 4 * link;
                        n0 = tex1Dfetch<float4>(sc.bvh_tex, t_); n1 = tex1Dfetch<float4>(sc.bvh_tex, t_ + 1);
                        n2 = tex1Dfetch<float4>(sc.bvh_tex, t_ + 2); n3 = tex1Dfetch<float4>(sc.bvh_tex, t_ + 3);
                    } else {
                        ldg8(nd, n0, n1); ldg8(nd + 2, n2, n3);
                    }
                    if (COUNT) wk.box += 2;
                    float a, b, l0, l1, r0, r1;
                    a = fmaf(n0.x, ix, -ox); b = fmaf(n0.y, ix, -ox); l0 = fminf(a, b); l1 = fmaxf(a, b);
                    a = fmaf(n0.z, iy, -oy); b = fmaf(n0.w, iy, -oy); l0 = fmaxf(l0, fminf(a, b)); l1 = fminf(l1, fmaxf(a, b));
                    a = fmaf(n2.x, iz, -oz); b = fmaf(n2.y, iz, -oz); l0 = fmaxf(l0, fminf(a, b)); l1 = fminf(l1, fmaxf(a, b));
                    a = fmaf(n1.x, ix, -ox); b = fmaf(n1.y, ix, -ox); r0 = fminf(a, b); r1 = fmaxf(a, b);
                    a = fmaf(n1.z, iy, -oy); b = fmaf(n1.w, iy, -oy); r0 = fmaxf(r0, fminf(a, b)); r1 = fminf(r1, fmaxf(a, b));
                    a = fmaf(n2.z, iz, -oz); b = fmaf(n2.w, iz, -oz); r0 = fmaxf(r0, fminf(a, b)); r1 = fminf(r1, fmaxf(a, b));
                    const float l0e = l0 * 0.999996f, r0e = r0 * 0.999996f;
                    const bool hl = (l0 <= l1 * 1.000004f) & (l1 >= 0.0f) & (l0e <= best.t);
                    const bool hr = (r0 <= r1 * 1.000004f) & (r1 >= 0.0f) & (r0e <= best.t);
                    const int ll = __float_as_int(n3.x), lr = __float_as_int(n3.y);
                    const bool right_first = hr & (!hl | (r0 < l0));          // both hit: the nearer entry first, left on ties
                    const int near_link = right_first ? lr : ll, far_link = right_first ? ll : lr;
                    if (hl & hr) WF_PUSH(far_link, right_first ? l0e : r0e);
                    if (WF_POSTPONE && (hl | hr) && near_link < 0 && pend == 0) { pend = near_link; st = WF_POP; }
                    else { link = near_link; st = (hl | hr) ? (near_link >= 0 ? WF_NODE : WF_LEAF) : WF_POP; }
                }
            }
        } else {
            // ------------------------------------------------------------ LEAF step
            // every lane with a leaf to test takes part: the postponed one first, then the one it is blocked on
            const int la = pend, lb = (st == WF_LEAF && link < 0) ? link : 0;
            if ((la | lb) < 0) {
                const int first_a = (~la) >> 3, cnt_a = la < 0 ? ((~la) & 7) + 1 : 0;
                const int first_b = (~lb) >> 3, cnt = cnt_a + (lb < 0 ? ((~lb) & 7) + 1 : 0);
                for (int k = 0; k < cnt; ++k) {
                    const int pos = k < cnt_a ? first_a + k : first_b + (k - cnt_a);
                    float t;
#if WF_TRI_RECORDS
                    float4 T0, T1, T2, T3;                           // one 64-byte record: two 256-bit loads, upload index included
                    ldg8(sc.srec + 4 * (size_t)pos, T0, T1);
                    ldg8(sc.srec + 4 * (size_t)pos + 2, T2, T3);
                    const bool acc_ = tri_accept<COUNT>(T0, T1, T2, P, D, t, wk);
                    if (acc_ & (t > 0.0f) & (t <= best.t)) {
                        const int tid = __float_as_int(T3.x);
                        if (t < best.t || tid < best.tri) { best.t = t; best.tri = tid; }
                    }
#else
                    const bool acc_ = tri_accept<COUNT>(ldg4(sc.sq0 + pos), ldg4(sc.sq1 + pos), ldg4(sc.sq2 + pos), P, D, t, wk);
                    if (acc_ & (t > 0.0f) & (t <= best.t)) {
                        const int tid = __ldg(sc.sorted_id + pos);
                        if (t < best.t || tid < best.tri) { best.t = t; best.tri = tid; }
                    }
#endif
                }
                pend = 0;
            }
            if (st == WF_LEAF) st = WF_POP;
        }
    }
#undef WF_PUSH
#undef WF_FINISH
#ifdef WF_DEBUG_STATES
    if (COUNT) { const bool l0_ = lane == 0; wk.box = l0_ ? dbgP : 0; wk.tri = 0; wk.tri1 = l0_ ? dbgN : 0; wk.tri2 = l0_ ? dbgL : 0; wk.tri3 = l0_ ? dbgF : 0; wk.samples = l0_ ? dbgI : 0; }
#endif
    wf_flush_counters(p, wk, COUNT);
}

// ---------------------------------------------------------------- host side: the iteration loop
// The frame's paths are dealt to WF_SUBS independent sub-wavefronts (blocks of 256 pixel ids, round-robin), each with its own
// queues and its own stream.  The persistent trace kernel of one sub-wavefront fills the machine; when its queue runs dry its
// blocks exit one by one, and the other sub-wavefront's shade / trace blocks move into the freed slots instead of the SMs
// idling until the last long ray of the launch is done (the tail is about one ray latency, ~40 us: 7 % of a 2 M-path
// iteration, 20 % of a 1 M-path one — the 8-GPU tile split).  Paths never move between sub-wavefronts and a pixel's samples
// stay sequential, so the image does not depend on the number of sub-wavefronts.
// which node format the trace kernel walks: 2 = quantised 4-ary (only when forced, ptb_set_quant_nodes(1): measured 3.37 vs 3.53 Grays/s on the
// 1 M-triangle scene and 3.67 vs 3.96 on the 104 k one — this kernel's node fetches are already split over both L1 data pipes and the decode
// costs it 24 more ALU instructions per node), 1 = float 4-ary, 0 = binary
static int wf_wide_mode(const SceneDev &sc) {
    if (sc.bvh4 == nullptr || sc.bvh_root < 0) return 0;
    return (sc.bvh4q != nullptr && sc.quant_wf) ? 2 : 1;
}
#define WF_SUBS_MAX 4
struct WfStreams { cudaStream_t s[WF_SUBS_MAX]; cudaEvent_t fork, join[WF_SUBS_MAX]; int *h_count; bool ok; };
static WfStreams &wf_streams() {
    static WfStreams w = [] {
        WfStreams t{};
        t.ok = true;
        for (int j = 1; j < WF_SUBS_MAX; ++j) {
            t.ok &= cudaStreamCreateWithFlags(&t.s[j], cudaStreamNonBlocking) == cudaSuccess;
            t.ok &= cudaEventCreateWithFlags(&t.join[j], cudaEventDisableTiming) == cudaSuccess;
        }
        t.ok &= cudaEventCreateWithFlags(&t.fork, cudaEventDisableTiming) == cudaSuccess;
        t.ok &= cudaMallocHost((void **)&t.h_count, WF_SUBS_MAX * sizeof(int)) == cudaSuccess;
        return t;
    }();
    return w;
}

int launch_wavefront(const RenderParams &p, cudaStream_t st) {
    const int n = p.W * p.H;
    int launches = 0;
    WfStreams &ws = wf_streams();
    const int want_subs = p.wf_subs > 0 ? p.wf_subs : 2;
    const int init_blocks = (n + 255) / 256;
    int subs = ws.ok ? (want_subs < 1 ? 1 : want_subs > WF_SUBS_MAX ? WF_SUBS_MAX : want_subs) : 1;
    if (subs > init_blocks) subs = init_blocks > 0 ? init_blocks : 1;
    const int sub_cap = ((init_blocks + subs - 1) / subs) * 256;          // queue slots per sub-wavefront (ptb_abi.cu sizes the queues for it)
    // every CUDA call of the loop is checked: the first failure ends the loop, -1 goes back to ptb_render (which reports the
    // CUDA error by name; a failed call leaves it in cudaGetLastError)
    bool failed = false;
#define WF_CK(call) do { if ((call) != cudaSuccess) failed = true; } while (0)
    WF_CK(cudaMemsetAsync(p.wf_ctrl, 0, 4 * WF_SUBS_MAX * sizeof(int), st));
    k_wf_init<<<init_blocks, 256, 0, st>>>(p, subs, sub_cap);
    WF_CK(cudaPeekAtLastError());
    if (failed) return -1;
    ++launches;
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // persistent grid: exactly as many blocks as are resident at once (a block that has to wait for a slot would start
    // after the queue is drained and add a tail); 54-64 registers and 16-20 KB of shared memory give 8-9 per SM
    static const int want_per_sm = getenv("PTB_WF_BLOCKS") ? atoi(getenv("PTB_WF_BLOCKS")) : 9;
    const int wide = wf_wide_mode(p.sc);
    int occ = 0;
    if (p.count_work) {
        if (wide == 2) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_wf_trace<true, 2>, PT_BLOCK, 0);
        else if (wide == 1) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_wf_trace<true, 1>, PT_BLOCK, 0);
        else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_wf_trace<true, 0>, PT_BLOCK, 0);
    } else {
        if (wide == 2) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_wf_trace<false, 2>, PT_BLOCK, 0);
        else if (wide == 1) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_wf_trace<false, 1>, PT_BLOCK, 0);
        else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_wf_trace<false, 0>, PT_BLOCK, 0);
    }
    const int per_sm = occ > 0 && occ < want_per_sm ? occ : want_per_sm;
    const int trace_blocks = sms * per_sm;
    static const int3 tune = make_int3(getenv("PTB_WF_FETCH_MIN") ? atoi(getenv("PTB_WF_FETCH_MIN")) : WF_FETCH_MIN,
                                       getenv("PTB_WF_LEAF_MIN") ? atoi(getenv("PTB_WF_LEAF_MIN")) : 4,
                                       getenv("PTB_WF_NODE_STEPS") ? atoi(getenv("PTB_WF_NODE_STEPS")) : PT_SCHED_NODE_STEPS);
    const long long max_iters = (long long)p.spp * (p.iterations > 0 ? p.iterations : 0);
    cudaStream_t sj[WF_SUBS_MAX];
    int host_count[WF_SUBS_MAX];
    bool live[WF_SUBS_MAX];
    sj[0] = st;
    // PTB_L2_PERSIST=1 (experiment): the wide-node array as a persisting L2 window on the trace streams, so that the path
    // state streaming through the shade kernel (80-110 B per path per iteration) does not evict the tree
    static const int l2_persist = getenv("PTB_L2_PERSIST") ? atoi(getenv("PTB_L2_PERSIST")) : 0;
    if (l2_persist && wide == 1 && p.sc.n_nodes4 > 0) {
        static int max_persist = -1, max_window = 0;
        if (max_persist < 0) {
            cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
            cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, dev);
            size_t want = (size_t)p.sc.n_nodes4 * 128;
            if (l2_persist == 1 || want > (size_t)max_persist) want = (size_t)max_persist;
            cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want);
            if (getenv("PTB_VERBOSE")) fprintf(stderr, "ptb: L2 persisting max %d B, window max %d B\n", max_persist, max_window);
        }
        size_t bytes = (size_t)p.sc.n_nodes4 * 128;
        if (bytes > (size_t)max_window) bytes = (size_t)max_window;
        cudaStreamAttrValue av;
        memset(&av, 0, sizeof av);
        av.accessPolicyWindow.base_ptr = const_cast<float4 *>(p.sc.bvh4);
        av.accessPolicyWindow.num_bytes = bytes;
        av.accessPolicyWindow.hitRatio = bytes <= (size_t)max_persist ? 1.0f : (float)max_persist / (float)bytes;
        av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        av.accessPolicyWindow.missProp = l2_persist == 1 ? cudaAccessPropertyStreaming : cudaAccessPropertyNormal;
        cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &av);
        for (int j = 1; j < subs; ++j) cudaStreamSetAttribute(ws.s[j], cudaStreamAttributeAccessPolicyWindow, &av);
    }
    if (subs > 1) WF_CK(cudaEventRecord(ws.fork, st));
    for (int j = 0; j < subs; ++j) {
        if (j > 0) { sj[j] = ws.s[j]; WF_CK(cudaStreamWaitEvent(sj[j], ws.fork, 0)); }
        host_count[j] = sub_cap;
        live[j] = true;
    }
    for (long long it = 0; it < max_iters && !failed; ++it) {
        const int a = (int)(it & 1), b = a ^ 1;
        bool any = false;
        for (int j = 0; j < subs; ++j) {
            if (!live[j]) continue;
            any = true;
            int *ctrl = p.wf_ctrl + 4 * j;
            int *qa = p.wf_queue[a] + (size_t)j * sub_cap, *qb = p.wf_queue[b] + (size_t)j * sub_cap;
            // next queue length and the fetch cursor start at zero
            WF_CK(cudaMemsetAsync(ctrl + b, 0, sizeof(int), sj[j]));
            WF_CK(cudaMemsetAsync(ctrl + 2, 0, sizeof(int), sj[j]));
            if (p.count_work) {
                if (wide == 2) k_wf_trace<true, 2><<<trace_blocks, PT_BLOCK, 0, sj[j]>>>(p, qa, ctrl + a, ctrl + 2, tune);
                else if (wide == 1) k_wf_trace<true, 1><<<trace_blocks, PT_BLOCK, 0, sj[j]>>>(p, qa, ctrl + a, ctrl + 2, tune);
                else k_wf_trace<true, 0><<<trace_blocks, PT_BLOCK, 0, sj[j]>>>(p, qa, ctrl + a, ctrl + 2, tune);
            } else {
                if (wide == 2) k_wf_trace<false, 2><<<trace_blocks, PT_BLOCK, 0, sj[j]>>>(p, qa, ctrl + a, ctrl + 2, tune);
                else if (wide == 1) k_wf_trace<false, 1><<<trace_blocks, PT_BLOCK, 0, sj[j]>>>(p, qa, ctrl + a, ctrl + 2, tune);
                else k_wf_trace<false, 0><<<trace_blocks, PT_BLOCK, 0, sj[j]>>>(p, qa, ctrl + a, ctrl + 2, tune);
            }
            WF_CK(cudaPeekAtLastError());
            k_wf_shade<<<(host_count[j] + 255) / 256, 256, 0, sj[j]>>>(p, qa, ctrl + a, qb, ctrl + b);
            WF_CK(cudaPeekAtLastError());
            launches += 2;
        }
        if (!any) break;
        if ((it & 15) == 15) {                     // every 16 iterations: learn the real queue lengths (shrinks the grids, ends early)
            for (int j = 0; j < subs; ++j)
                if (live[j]) WF_CK(cudaMemcpyAsync(ws.ok ? ws.h_count + j : host_count + j, p.wf_ctrl + 4 * j + b, sizeof(int), cudaMemcpyDeviceToHost, sj[j]));
            for (int j = 0; j < subs; ++j) {
                if (!live[j]) continue;
                WF_CK(cudaStreamSynchronize(sj[j]));
                if (failed) break;
                if (ws.ok) host_count[j] = ws.h_count[j];
                if (host_count[j] == 0) live[j] = false;
            }
        }
    }
    for (int j = 1; j < subs; ++j) {               // join: the caller's stream continues after every sub-wavefront
        WF_CK(cudaEventRecord(ws.join[j], sj[j]));
        WF_CK(cudaStreamWaitEvent(st, ws.join[j], 0));
    }
#undef WF_CK
    return failed ? -1 : launches;
}
