// pt_pool.cuh — the LBVH path as ONE persistent kernel whose warps regroup their rays by traversal state
// (included inside namespace PT_NS after pt_wavefront.cuh, whose path-state helpers it shares).
//
// Why (profiles/r01_wf_trace_v9.txt): in k_wf_trace a lane owns one ray for the ray's whole traversal, so the lanes of a
// warp sit in different states — 13.8 of 32 lanes active per instruction (wide-node step 17 lanes, leaf step 6, pop 5) —
// and the host drives spp x depth iterations of trace + shade launches whose tails and fixed costs eat 24 % of an
// 8-way tile split.  Here:
//   * a WARP owns a pool of POOL_NS ray slots (default 48) whose traversal state lives in SHARED memory, slot-indexed:
//     origin, direction, reciprocal direction, best hit, current link, stack pointer, the first POOL_STACK stack entries
//     (deeper ones in a global overflow area), pixel id and a state tag;
//   * every warp step takes a census of the tags (six ballots), picks ONE phase — NODE (wide-node slab tests incl. the
//     pop that precedes them), LEAF (triangle tests) or SHADE (hit -> material -> bounce / next sample / next pixel, the
//     outlier scan of the new ray) — and hands the first 32 slots waiting for that phase to its 32 lanes (rank by
//     prefix popcount through a 32-entry shared array).  Any lane can serve any slot, so a node step runs at 32 lanes as
//     long as 32 of the 48 slots want one, instead of at however many of a warp's 32 private rays happen to;
//   * a slot keeps its PIXEL until all samples of the pixel are done (samples of a pixel are sequential: same RNG order
//     as the reference), then takes the next pixel from a global cursor.  There are no ray queues, no global ray
//     records, no iterations and no host loop: ptb_render is one memset + one launch, and nothing waits for a tail
//     except the very end of the render;
//   * path state (throughputs, sample colour, accumulator, RNG: 6 float4) lives in global memory indexed by
//     (warp, slot), touched only by the SHADE phase through L2 (ld.cg / st.cg: another lane of the warp may serve the
//     slot next time).
// MEASURED AND DROPPED (commit ac1d6e8, profiles/r02_pool_shards.txt): cutting every pixel's samples into chunks handed out chunk-major,
// each chunk starting from the (RNG state, accumulator) its predecessor published in rnds[] / colors[], so that a GPU that owns only ~5
// pixels per slot does not end with most slots idle during the last round of pixels: bit-identical, but 2570 vs 2733 Mrays/s on rank 0's
// share of an 8-way tile split of the 4K frame and 3135 vs 3213 on the full 1080p frame — the extra per-item state reads, fences and
// registers cost more than the shorter tail returns.
// The traversal is the wavefront kernel's: 4-ary nodes, near/far planes picked by direction sign, 64-byte triangle
// records, entry-distance culling at pop, closest hit = min (t, upload index).  Bit-identical to the other pipelines in
// the parity flavour (tests/test_gpu_pool.py).
#ifndef POOL_NS
#define POOL_NS 48            // ray slots per warp
#endif
#ifndef POOL_STACK
#define POOL_STACK 8          // stack entries per slot kept in shared memory
#endif
#ifndef POOL_MINB
#define POOL_MINB 7           // resident blocks per SM the register allocation is held to
#endif
#ifndef POOL_QN_FETCH
#define POOL_QN_FETCH 2       // quantised nodes through the texture pipe (see the node step)
#endif
#define POOL_WARPS 4
#define POOL_DEEP (PT_LBVH_STACK - POOL_STACK)
enum { PL_DEAD = 0, PL_NODE = 1, PL_POP = 2, PL_LEAF = 3, PL_SHADE = 4, PL_EMPTY = 5 };
enum { PW_PX = 0, PW_PY, PW_PZ, PW_DX, PW_DY, PW_DZ, PW_IX, PW_IY, PW_IZ, PW_BT, PW_BTRI, PW_LINK, PW_SP, PW_PIX, PW_TAG, PW_FIELDS };
#define POOL_WORDS (PW_FIELDS * POOL_NS + 2 * POOL_STACK * POOL_NS + 32)       // 32-bit words of shared memory per warp

struct PoolArgs {
    int *cursor;                 // next item (owned pixel / ray of the list) to hand out; zeroed before the launch
    int n_items;
    int2 *deep;                  // stack overflow, POOL_DEEP entries per slot
    float4 *state;               // WF_PLANES planes of total_slots float4
    int total_slots;
    int leaf_min, shade_min, node_steps;
    const RefRay *rays;          // closest-hit mode: the ray list and its result pairs
    float2 *out;
};

// reciprocal direction for the slab tests: bounded to 1e25 so that the quantised-node decode (128 * step / d, up to 2^20 / d) cannot
// overflow; a component that small means "parallel to the slabs of this axis" for every scene a float can describe
PT_DEV float pool_rcp(float d) { return fabsf(d) > 1e-25f ? 1.0f / d : copysignf(1e25f, d); }

// Path state of a slot: 24 words as THREE 32-byte chunks (planes of total_slots chunks), each one 256-bit access: 3 loads + 3 stores
// per shaded ray on the LSU pipe instead of 6 + 6 sixteen-byte ones.  Through L2 (.cg): another lane of the warp may serve the slot next.
// MEASURED AND DROPPED (profiles/r02_pool_state_split.txt): the wavefront's hot / cold split (two chunks every step, the third — running mean,
// sample index, sample radiance — only when a sample ends or an emitter is hit) is 2.5 % SLOWER here (3.76 vs 3.86 Grays/s on the 104 k scene,
// 2.94 vs 3.00 on a 1/8 share of the 1 M scene), with the third load conditional or not: the state lives in L2 next to the warp, its bytes
// are not what the shade phase waits for, and the extra predicates and the dirty flag cost issue slots in the kernel's longest phase.
PT_DEV void pool_st8(float4 *p, float a, float b, float c, float d, float e, float f, float g, float h) {
    asm volatile("st.global.cg.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d), "f"(e), "f"(f), "f"(g), "f"(h) : "memory");
}
PT_DEV void pool_ld8(const float4 *p, float4 &a, float4 &b) {
    asm volatile("ld.global.cg.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p) : "memory");
}
PT_DEV void pool_store_state(const PoolArgs &a, int g, const WfPath &w) {
    float4 *st = a.state + 2 * (size_t)g;
    const size_t n = 2 * (size_t)a.total_slots;
    pool_st8(st, w.ps.fL.x, w.ps.fL.y, w.ps.fL.z, w.ps.fB.x, w.ps.fB.y, w.ps.fB.z, w.ps.fS.x, w.ps.fS.y);
    pool_st8(st + n, w.ps.fS.z, w.ps.fR.x, w.ps.fR.y, w.ps.fR.z, w.ps.color.x, w.ps.color.y, w.ps.color.z, __uint_as_float(w.rng));
    pool_st8(st + 2 * n, w.acc.x, w.acc.y, w.acc.z, __int_as_float(w.s), __int_as_float(w.depth), __int_as_float(w.ps.in ? 1 : 0), 0.0f, 0.0f);
}
PT_DEV void pool_load_state(const PoolArgs &a, int g, WfPath &w) {
    const float4 *st = a.state + 2 * (size_t)g;
    const size_t n = 2 * (size_t)a.total_slots;
    float4 A, B, C, D, E, F;
    pool_ld8(st, A, B); pool_ld8(st + n, C, D); pool_ld8(st + 2 * n, E, F);
    w.ps.fL = v3(A.x, A.y, A.z); w.ps.fB = v3(A.w, B.x, B.y); w.ps.fS = v3(B.z, B.w, C.x); w.ps.fR = v3(C.y, C.z, C.w);
    w.ps.color = v3(D.x, D.y, D.z); w.rng = __float_as_uint(D.w);
    w.acc = v3(E.x, E.y, E.z); w.s = __float_as_int(E.w);
    w.depth = __float_as_int(F.x); w.ps.in = __float_as_int(F.y) != 0;
}
// k-th pixel this rank owns under ptb_set_pixel_shard (bands of `group` consecutive ids dealt round-robin)
PT_DEV int pool_item_pixel(const RenderParams &p, int k) {
    if (p.shard_n <= 1) return k;
    const int band = k / p.shard_group, off = k - band * p.shard_group;
    return (band * p.shard_n + p.shard_rank) * p.shard_group + off;
}

// MODE 0: render (pixels, path state).  MODE 1: closest hit of a ray list (ptb_closest_hit), same traversal phases.
// QN: the wide nodes are read in their 64-byte quantised form (lbvh.cu k_emit4q): two 256-bit loads per node instead of five fetches.
template <bool COUNT, int MODE, bool QN>
__global__ void __launch_bounds__(32 * POOL_WARPS, POOL_MINB) k_pool(const RenderParams p, const PoolArgs a) {
    extern __shared__ int pool_smem[];
    __shared__ float4 s_big[WF_BIG_SMEM];
    const SceneDev &sc = p.sc;
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gwarp = blockIdx.x * POOL_WARPS + warp;
    const unsigned lt = (1u << lane) - 1u;
    int *wi = pool_smem + warp * POOL_WORDS;
    float *wf = reinterpret_cast<float *>(wi);
    int2 *wstack = reinterpret_cast<int2 *>(wi + PW_FIELDS * POOL_NS);          // [k * POOL_NS + slot]
    int *sel = wi + PW_FIELDS * POOL_NS + 2 * POOL_STACK * POOL_NS;
    int2 *deep = a.deep + (size_t)gwarp * POOL_NS * POOL_DEEP;                   // [slot * POOL_DEEP + k]
#define PF(f, s) wf[(f) * POOL_NS + (s)]
#define PI(f, s) wi[(f) * POOL_NS + (s)]
    const int n_big_pairs = MODE == 0 ? wf_stage_big(p, s_big) : 0;
    Work wk = {0, 0, 0, 0, 0, 0, 0};
#ifdef POOL_DEBUG_STATES   // throw-away instrumentation (scripts/pool_states.py): warp steps and served slots per phase, through the work counters
    unsigned dbg_steps[3] = {0, 0, 0}, dbg_lanes[3] = {0, 0, 0};
#endif
    bool exhausted = a.n_items <= 0;                  // warp-uniform: the cursor has handed out every item
    const V3 eye = v3(p.cam.eye.x, p.cam.eye.y, p.cam.eye.z);
    const int s_end = p.sample0 + p.spp;
    for (int s = lane; s < POOL_NS; s += 32) PI(PW_TAG, s) = PL_EMPTY;
    __syncwarp();
    for (;;) {
        // ------------------------------------------------------------ census of the pool, one phase for the whole warp
        const int t0 = PI(PW_TAG, lane);
        const int t1 = lane + 32 < POOL_NS ? PI(PW_TAG, (lane + 32) % POOL_NS) : PL_DEAD;
        const bool isN0 = t0 == PL_NODE || t0 == PL_POP, isN1 = t1 == PL_NODE || t1 == PL_POP;
        const bool isL0 = t0 == PL_LEAF, isL1 = t1 == PL_LEAF, isS0 = t0 >= PL_SHADE, isS1 = t1 >= PL_SHADE;
        const unsigned counts = __reduce_add_sync(full, (unsigned)(isN0 + isN1) | ((unsigned)(isL0 + isL1) << 8) | ((unsigned)(isS0 + isS1) << 16));   // one REDUX
        const int cN = counts & 0xff, cL = (counts >> 8) & 0xff, cS = counts >> 16;
        if (counts == 0u) break;                      // every slot is dead: no pixels left
        int phase;                                    // 0 node, 1 leaf, 2 shade
        if (cN >= 32) phase = 0;
        else if (cL >= a.leaf_min) phase = 1;
        else if (cS >= a.shade_min) phase = 2;
        else phase = (cN >= cL && cN >= cS) ? 0 : (cL >= cS ? 1 : 2);
        const unsigned m0 = __ballot_sync(full, phase == 0 ? isN0 : (phase == 1 ? isL0 : isS0));
        const unsigned m1 = __ballot_sync(full, phase == 0 ? isN1 : (phase == 1 ? isL1 : isS1));
        const int c0 = __popc(m0);
        if ((m0 >> lane) & 1u) sel[__popc(m0 & lt)] = lane;
        if ((m1 >> lane) & 1u) { const int r = c0 + __popc(m1 & lt); if (r < 32) sel[r] = lane + 32; }
        __syncwarp();
        const int cnt = min(c0 + __popc(m1), 32);
        const bool active = lane < cnt;
        const int slot = active ? sel[lane] : 0;
#ifdef POOL_DEBUG_STATES
        dbg_steps[phase] += 1; dbg_lanes[phase] += cnt;
#endif
        if (phase == 0) {
            // -------------------------------------------------------- NODE: pop (if the slot needs one), then the wide node's four slab tests
            int tg = PL_DEAD, link = 0, sp = 0;
            float bt = 0.0f, ix = 0.0f, iy = 0.0f, iz = 0.0f, ox = 0.0f, oy = 0.0f, oz = 0.0f;
            if (active) {
                tg = PI(PW_TAG, slot); link = PI(PW_LINK, slot); sp = PI(PW_SP, slot); bt = PF(PW_BT, slot);
                ix = PF(PW_IX, slot); iy = PF(PW_IY, slot); iz = PF(PW_IZ, slot);
                ox = PF(PW_PX, slot) * ix; oy = PF(PW_PY, slot) * iy; oz = PF(PW_PZ, slot) * iz;
            }
            const int nx = ix < 0.0f ? 1 : 0, ny = iy < 0.0f ? 1 : 0, nz = iz < 0.0f ? 1 : 0;     // row of the NEAR plane per axis
#define PL_PUSH(v, d) do { const int2 e_ = make_int2((v), __float_as_int(d)); if (sp < POOL_STACK) wstack[sp * POOL_NS + slot] = e_; \
                           else if (sp < PT_LBVH_STACK) __stcg(deep + slot * POOL_DEEP + (sp - POOL_STACK), e_); ++sp; } while (0)
#pragma unroll 1
            for (int k = 0; k < a.node_steps; ++k) {
                if (k > 0 && !__any_sync(full, tg == PL_NODE || tg == PL_POP)) break;
                if (tg == PL_POP) {               // one attempt per step: an entry is culled here only if a leaf step shrank the best hit since it was pushed
                    if (sp == 0) tg = PL_SHADE;   // traversal finished: the best hit is in the slot
                    else {
                        --sp;
                        const int2 e = sp < POOL_STACK ? wstack[sp * POOL_NS + slot] : __ldcg(deep + slot * POOL_DEEP + (sp - POOL_STACK));
                        if (__int_as_float(e.y) <= bt) { link = e.x; tg = e.x >= 0 ? PL_NODE : PL_LEAF; }
                    }
                }
                if (tg == PL_NODE && QN) {
                    // Quantised node: child c's slab on axis a is [p_a + qlo * step_a, p_a + qhi * step_a].  A byte q becomes the float
                    // f = 256 + q / 128 with ONE byte permute (q into byte 1 of 0x43800000), so t = (p + q step - O) / d =
                    // fma(f, A, C) with A = 128 step / d and C = (p - O) / d - 256 A: one permute + one FMA per plane.  The only
                    // rounding beyond the float-box test's is that of C: < 2^-15 |A| = 2^-8 step, inside the whole step the
                    // builder left on every side of every child box.
                    const uint4 *nq = reinterpret_cast<const uint4 *>(sc.bvh4q) + 4 * (size_t)link;
                    unsigned w0, w1, w2, w3, w4, w5, w6, w7, w8, w9, w10, w11, k0u, k1u, k2u, k3u;
                    // POOL_QN_FETCH: which L1 data pipe carries the node (the LSU pipe also carries the pool's shared-memory traffic and
                    // is the busier one, profiles/r02_pool_v1.txt): 0 = two 256-bit LSU loads, 1 = first half LSU + second half two texture
                    // fetches, 2 = four texture fetches
                    if (POOL_QN_FETCH == 2 && sc.bvh4q_tex) {
                        const uint4 a0 = tex1Dfetch<uint4>(sc.bvh4q_tex, 4 * link), a1 = tex1Dfetch<uint4>(sc.bvh4q_tex, 4 * link + 1);
                        const uint4 a2 = tex1Dfetch<uint4>(sc.bvh4q_tex, 4 * link + 2), a3 = tex1Dfetch<uint4>(sc.bvh4q_tex, 4 * link + 3);
                        w0 = a0.x; w1 = a0.y; w2 = a0.z; w3 = a0.w; w4 = a1.x; w5 = a1.y; w6 = a1.z; w7 = a1.w;
                        w8 = a2.x; w9 = a2.y; w10 = a2.z; w11 = a2.w; k0u = a3.x; k1u = a3.y; k2u = a3.z; k3u = a3.w;
                    } else {
                        asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                     : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3), "=r"(w4), "=r"(w5), "=r"(w6), "=r"(w7) : "l"(nq));
                        if (POOL_QN_FETCH == 1 && sc.bvh4q_tex) {
                            const uint4 a2 = tex1Dfetch<uint4>(sc.bvh4q_tex, 4 * link + 2), a3 = tex1Dfetch<uint4>(sc.bvh4q_tex, 4 * link + 3);
                            w8 = a2.x; w9 = a2.y; w10 = a2.z; w11 = a2.w; k0u = a3.x; k1u = a3.y; k2u = a3.z; k3u = a3.w;
                        } else {
                            asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                         : "=r"(w8), "=r"(w9), "=r"(w10), "=r"(w11), "=r"(k0u), "=r"(k1u), "=r"(k2u), "=r"(k3u) : "l"(nq + 2));
                        }
                    }
                    (void)w10; (void)w11;
                    if (COUNT) wk.box += 4;
                    const float Ax = __uint_as_float((w3 & 0xffu) << 23) * ix, Ay = __uint_as_float(((w3 >> 8) & 0xffu) << 23) * iy,
                                Az = __uint_as_float(((w3 >> 16) & 0xffu) << 23) * iz;
                    const float Cx = fmaf(-256.0f, Ax, fmaf(__uint_as_float(w0), ix, -ox)), Cy = fmaf(-256.0f, Ay, fmaf(__uint_as_float(w1), iy, -oy)),
                                Cz = fmaf(-256.0f, Az, fmaf(__uint_as_float(w2), iz, -oz));
                    const unsigned nxw = nx ? w5 : w4, fxw = nx ? w4 : w5, nyw = ny ? w7 : w6, fyw = ny ? w6 : w7, nzw = nz ? w9 : w8, fzw = nz ? w8 : w9;
                    float d0, d1, d2, d3;
#define PL_QF(word, sel) __uint_as_float(__byte_perm((word), 0x43800000u, (sel)))
#define PL_QSLAB(dk, sel) do { float lo_ = fmaxf(fmaxf(fmaf(PL_QF(nxw, sel), Ax, Cx), fmaf(PL_QF(nyw, sel), Ay, Cy)), fmaf(PL_QF(nzw, sel), Az, Cz));   \
                               float hi_ = fminf(fminf(fmaf(PL_QF(fxw, sel), Ax, Cx), fmaf(PL_QF(fyw, sel), Ay, Cy)), fmaf(PL_QF(fzw, sel), Az, Cz));   \
                               lo_ = fmaxf(lo_ * 0.999996f, 0.0f); hi_ = fminf(hi_ * 1.000004f, bt);                                                    \
                               dk = lo_ <= hi_ ? lo_ : PT_NO_HIT; } while (0)
                    PL_QSLAB(d0, 0x7604); PL_QSLAB(d1, 0x7614); PL_QSLAB(d2, 0x7624); PL_QSLAB(d3, 0x7634);
#undef PL_QSLAB
#undef PL_QF
                    int k0 = (int)k0u, k1 = (int)k1u, k2 = (int)k2u, k3 = (int)k3u;
#define PL_CE(da, ka, db, kb) do { const bool sw_ = db < da; const float td_ = sw_ ? db : da; const int tk_ = sw_ ? kb : ka; \
                                   db = sw_ ? da : db; kb = sw_ ? ka : kb; da = td_; ka = tk_; } while (0)
                    PL_CE(d0, k0, d1, k1); PL_CE(d2, k2, d3, k3); PL_CE(d0, k0, d2, k2); PL_CE(d1, k1, d3, k3); PL_CE(d1, k1, d2, k2);
#undef PL_CE
                    if (d3 < PT_NO_HIT) PL_PUSH(k3, d3);
                    if (d2 < PT_NO_HIT) PL_PUSH(k2, d2);
                    if (d1 < PT_NO_HIT) PL_PUSH(k1, d1);
                    link = k0;
                    tg = d0 < PT_NO_HIT ? (k0 >= 0 ? PL_NODE : PL_LEAF) : PL_POP;
                }
                if (tg == PL_NODE && !QN) {
                    // see k_wf_trace (pt_wavefront.cuh) for the layout and the conservative slab test; same arithmetic
                    const float4 *nd = sc.bvh4 + 8 * (size_t)link;
                    float4 Xn, Xf, Yn, Yf, Zn, Zf, LK;
                    if (sc.bvh4_tex) {
                        const int t_ = 8 * link;
                        Xn = tex1Dfetch<float4>(sc.bvh4_tex, t_ + nx); Xf = tex1Dfetch<float4>(sc.bvh4_tex, t_ + (nx ^ 1));
                        { float4 A0, A1; ldg8(nd + 2, A0, A1); Yn = ny ? A1 : A0; Yf = ny ? A0 : A1; }
                        { float4 A0, A1; ldg8(nd + 4, A0, A1); Zn = nz ? A1 : A0; Zf = nz ? A0 : A1; }
                        LK = tex1Dfetch<float4>(sc.bvh4_tex, t_ + 6);
                    } else {
                        Xn = ldg4(nd + nx); Xf = ldg4(nd + (nx ^ 1)); Yn = ldg4(nd + 2 + ny); Yf = ldg4(nd + 2 + (ny ^ 1));
                        Zn = ldg4(nd + 4 + nz); Zf = ldg4(nd + 4 + (nz ^ 1)); LK = ldg4(nd + 6);
                    }
                    if (COUNT) wk.box += 4;
                    float d0, d1, d2, d3;
#define PL_SLAB(dk, c) do { float lo_ = fmaxf(fmaxf(fmaf(Xn.c, ix, -ox), fmaf(Yn.c, iy, -oy)), fmaf(Zn.c, iz, -oz));                 \
                            float hi_ = fminf(fminf(fmaf(Xf.c, ix, -ox), fmaf(Yf.c, iy, -oy)), fmaf(Zf.c, iz, -oz));                 \
                            lo_ = fmaxf(lo_ * 0.999996f, 0.0f); hi_ = fminf(hi_ * 1.000004f, bt);                                    \
                            dk = lo_ <= hi_ ? lo_ : PT_NO_HIT; } while (0)
                    PL_SLAB(d0, x); PL_SLAB(d1, y); PL_SLAB(d2, z); PL_SLAB(d3, w);
#undef PL_SLAB
                    int k0 = __float_as_int(LK.x), k1 = __float_as_int(LK.y), k2 = __float_as_int(LK.z), k3 = __float_as_int(LK.w);
#define PL_CE(da, ka, db, kb) do { const bool sw_ = db < da; const float td_ = sw_ ? db : da; const int tk_ = sw_ ? kb : ka; \
                                   db = sw_ ? da : db; kb = sw_ ? ka : kb; da = td_; ka = tk_; } while (0)
                    PL_CE(d0, k0, d1, k1); PL_CE(d2, k2, d3, k3); PL_CE(d0, k0, d2, k2); PL_CE(d1, k1, d3, k3); PL_CE(d1, k1, d2, k2);
#undef PL_CE
                    if (d3 < PT_NO_HIT) PL_PUSH(k3, d3);
                    if (d2 < PT_NO_HIT) PL_PUSH(k2, d2);
                    if (d1 < PT_NO_HIT) PL_PUSH(k1, d1);
                    link = k0;
                    tg = d0 < PT_NO_HIT ? (k0 >= 0 ? PL_NODE : PL_LEAF) : PL_POP;
                }
            }
#undef PL_PUSH
            if (active) { PI(PW_TAG, slot) = tg; PI(PW_LINK, slot) = link; PI(PW_SP, slot) = sp; }
        } else if (phase == 1) {
            // -------------------------------------------------------- LEAF: the slot's <= 8 triangles, 64-byte records (two 256-bit loads each)
            if (active) {
                const int enc = ~PI(PW_LINK, slot);
                const int first = enc >> 3, count = (enc & 7) + 1;
                const V3 P = v3(PF(PW_PX, slot), PF(PW_PY, slot), PF(PW_PZ, slot)), D = v3(PF(PW_DX, slot), PF(PW_DY, slot), PF(PW_DZ, slot));
                Closest best; best.t = PF(PW_BT, slot); best.tri = PI(PW_BTRI, slot);
                for (int k = 0; k < count; ++k) {
                    const int pos = first + k;
                    float4 T0, T1, T2, T3;
                    ldg8(sc.srec + 4 * (size_t)pos, T0, T1);
                    ldg8(sc.srec + 4 * (size_t)pos + 2, T2, T3);
                    float t;
                    const bool acc_ = tri_accept<COUNT>(T0, T1, T2, P, D, t, wk);
                    if (acc_ & (t > 0.0f) & (t <= best.t)) {
                        const int tid = __float_as_int(T3.x);
                        if (t < best.t || tid < best.tri) { best.t = t; best.tri = tid; }
                    }
                }
                PF(PW_BT, slot) = best.t; PI(PW_BTRI, slot) = best.tri; PI(PW_TAG, slot) = PL_POP;
            }
        } else {
            // -------------------------------------------------------- SHADE: finished rays -> next ray of the path / next sample / next pixel
            const int tg = active ? PI(PW_TAG, slot) : PL_DEAD;
            const int gslot = gwarp * POOL_NS + slot;
            int pix = -1;
            WfPath w;
            w.rng = 1u; w.acc = v3(0.0f, 0.0f, 0.0f); w.s = s_end; w.depth = 0;
            path_reset(w.ps);
            V3 rayP = v3(0.0f, 0.0f, 0.0f), rayD = v3(0.0f, 0.0f, 1.0f);
            bool has = false;                         // the slot leaves this phase with a ray to trace
            bool want_sample = false;                 // ... needs the next sample of its pixel started (or the pixel finished)
            bool want_pixel = active && tg == PL_EMPTY;
            if (tg == PL_SHADE) {
                pix = PI(PW_PIX, slot);
                const float t = PF(PW_BT, slot);
                const int tri = PI(PW_BTRI, slot);
                if (MODE == 1) {
                    const bool hit = t < PT_NO_HIT;
                    a.out[pix] = make_float2(hit ? t : -1.0f, __int_as_float(hit ? tri : -1));
                    want_pixel = true;
                } else {
                    pool_load_state(a, gslot, w);
                    rayP = v3(PF(PW_PX, slot), PF(PW_PY, slot), PF(PW_PZ, slot)); rayD = v3(PF(PW_DX, slot), PF(PW_DY, slot), PF(PW_DZ, slot));
                    wk.rays++;
                    if (t < PT_NO_HIT) {
                        const V3 hitP = rayP + rayD * t;                         // prog.cl:107
                        const V3 hitN = tri_normal(ldg4(sc.q0 + tri), ldg4(sc.q1 + tri), ldg4(sc.q2 + tri));
                        const Mat m = load_material(sc.mats, (int)__ldg(sc.mati + tri));
                        shade_hit(m, hitP, hitN, rayP, rayD, w.rng, w.ps, eye, p.iterations);
                        has = ++w.depth < p.iterations;
                    }                                                            // miss: break, prog.cl:367-376
                    if (!has) {                                                  // sample finished, prog.cl:379
                        if (p.sum_mode) w.acc = w.acc + w.ps.color; else w.acc = pt_mean_step(w.acc, w.ps.color, w.s);
                        ++w.s;
                        wk.samples++;
                        want_sample = true;
                    }
                }
            }
            // Rounds of "take the next pixel of the cursor" (one atomic per warp and round) and "start the next sample" until every
            // served slot has a ray or the cursor has run dry.  One call site each for the camera ray and the end-of-pixel stores.
            for (;;) {
                const unsigned need = __ballot_sync(full, want_pixel && !exhausted);
                if (need != 0u) {
                    const int leader = __ffs(need) - 1;
                    int base = 0;
                    if (lane == leader) base = atomicAdd(a.cursor, __popc(need));
                    base = __shfl_sync(full, base, leader);
                    if (base + __popc(need) >= a.n_items) exhausted = true;
                    const int k = base + __popc(need & lt);
                    if (((need >> lane) & 1u) && k < a.n_items) {
                        want_pixel = false;
                        if (MODE == 1) {
                            const float4 *r = reinterpret_cast<const float4 *>(a.rays + k);
                            const float4 r0 = ldg4(r), r1 = ldg4(r + 1);
                            rayP = v3(r0.x, r0.y, r0.z); rayD = v3(r1.x, r1.y, r1.z);
                            pix = k;
                            has = true;
                        } else {
                            pix = pool_item_pixel(p, k);
                            w.rng = (uint32_t)p.rnds[pix];
                            w.acc = (p.sample0 != 0 || p.sum_mode) ? v3(p.colors[pix]) : v3(0.0f, 0.0f, 0.0f);
                            w.s = p.sample0;
                            want_sample = true;
                        }
                    }
                }
                if (MODE == 1) { if (need == 0u) break; continue; }
                if (!__any_sync(full, want_sample)) { if (need == 0u) break; continue; }
                if (want_sample) {
                    want_sample = false;
                    if (w.s < s_end) {                                           // gen_ray, prog.cl:384-389
                        const int camX = max((int)p.cam.XM, 1);
                        const float r1 = pt_rand(w.rng), r2 = pt_rand(w.rng);
                        camera_ray(p, pix % camX, pix / camX, r1, r2, rayP, rayD);
                        path_reset(w.ps);
                        w.depth = 0;
                        if (p.iterations > 0) has = true;
                        else {                                                   // no bounce loop: the sample is black, only the jitter draws happen
                            if (p.sum_mode) w.acc = w.acc + w.ps.color; else w.acc = pt_mean_step(w.acc, w.ps.color, w.s);
                            ++w.s;
                            wk.samples++;
                            want_sample = true;
                        }
                    } else {
                        wf_finish_pixel(p, pix, w);
                        pix = -1;
                        want_pixel = true;
                    }
                }
            }
            if (active) {
                if (has) {
                    Closest best;
                    if (MODE == 1) { best.t = PT_NO_HIT; best.tri = 0x7fffffff; lbvh_big_scan<false>(sc, rayP, rayD, best, wk); }
                    else wf_big_scan(p, s_big, n_big_pairs, rayP, rayD, best, wk);
                    PF(PW_PX, slot) = rayP.x; PF(PW_PY, slot) = rayP.y; PF(PW_PZ, slot) = rayP.z;
                    PF(PW_DX, slot) = rayD.x; PF(PW_DY, slot) = rayD.y; PF(PW_DZ, slot) = rayD.z;
                    PF(PW_IX, slot) = pool_rcp(rayD.x); PF(PW_IY, slot) = pool_rcp(rayD.y); PF(PW_IZ, slot) = pool_rcp(rayD.z);
                    PF(PW_BT, slot) = best.t; PI(PW_BTRI, slot) = best.tri;
                    PI(PW_LINK, slot) = sc.bvh_root; PI(PW_SP, slot) = 0; PI(PW_PIX, slot) = pix;
                    PI(PW_TAG, slot) = sc.bvh_root >= 0 ? PL_NODE : PL_LEAF;
                    if (MODE == 0) pool_store_state(a, gslot, w);
                } else {
                    PI(PW_TAG, slot) = PL_DEAD;
                }
            }
        }
        __syncwarp();
    }
#undef PF
#undef PI
#ifdef POOL_DEBUG_STATES
    if (COUNT) { const bool l0_ = lane == 0; wk.box = l0_ ? dbg_steps[0] : 0; wk.tri = l0_ ? dbg_lanes[0] : 0; wk.tri1 = l0_ ? dbg_steps[1] : 0;
                 wk.tri2 = l0_ ? dbg_lanes[1] : 0; wk.tri3 = l0_ ? dbg_steps[2] : 0; wk.samples = l0_ ? dbg_lanes[2] : 0; }
#endif
    wf_flush_counters(p, wk, COUNT);
}

// ---------------------------------------------------------------- host side
struct PoolPlan { int blocks; int total_slots; size_t smem; size_t scratch_bytes; };
template <bool COUNT, int MODE, bool QN> static PoolPlan pool_plan_t() {
    static PoolPlan plans[16];
    static bool have[16] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    dev &= 15;
    if (have[dev]) return plans[dev];            // function attributes are per device: set them once on each
    PoolPlan pl{};
    int sms = 148, occ = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    pl.smem = (size_t)POOL_WARPS * POOL_WORDS * sizeof(int);
    if (pl.smem > 48 * 1024) cudaFuncSetAttribute(k_pool<COUNT, MODE, QN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    cudaFuncSetAttribute(k_pool<COUNT, MODE, QN>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_pool<COUNT, MODE, QN>, 32 * POOL_WARPS, pl.smem);
    static const int want = getenv("PTB_POOL_BLOCKS") ? atoi(getenv("PTB_POOL_BLOCKS")) : 0;
    if (want > 0 && want < occ) occ = want;
    if (occ < 1) occ = 1;
    // the carve-out the resident blocks need (+ 1 KB per block the runtime reserves) and nothing more: the rest stays L1 for nodes and triangles
    int pct = (int)((occ * (pl.smem + sizeof(float4) * WF_BIG_SMEM + 1024) * 100 + 228 * 1024 - 1) / (228 * 1024));
    if (pct > 100) pct = 100;
    cudaFuncSetAttribute(k_pool<COUNT, MODE, QN>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
    pl.blocks = sms * occ;
    pl.total_slots = pl.blocks * POOL_WARPS * POOL_NS;
    pl.scratch_bytes = 256 + (size_t)pl.total_slots * POOL_DEEP * sizeof(int2) + 256 + (size_t)pl.total_slots * WF_PLANES * sizeof(float4) + 256;
    if (getenv("PTB_VERBOSE")) fprintf(stderr, "ptb: k_pool<%d,%d,%d> on device %d: %d blocks/SM, %zu B smem/block, carve-out %d %%, %d slots, scratch %.1f MB\n",
                                       (int)COUNT, MODE, (int)QN, dev, occ, pl.smem, pct, pl.total_slots, pl.scratch_bytes / 1048576.0);
    plans[dev] = pl;
    have[dev] = true;
    return pl;
}
// bytes of scratch the pool kernels need (the C-ABI allocates them once)
size_t pool_scratch_bytes() {
    size_t m = pool_plan_t<false, 0, true>().scratch_bytes;
    const size_t o[5] = {pool_plan_t<true, 0, true>().scratch_bytes, pool_plan_t<false, 1, true>().scratch_bytes, pool_plan_t<false, 0, false>().scratch_bytes,
                         pool_plan_t<true, 0, false>().scratch_bytes, pool_plan_t<false, 1, false>().scratch_bytes};
    for (size_t v : o) if (v > m) m = v;
    return m;
}
static PoolArgs pool_args(const PoolPlan &pl, void *scratch, int n_items) {
    PoolArgs a{};
    char *b = (char *)scratch;
    a.cursor = (int *)b; b += 256;
    a.deep = (int2 *)b; b += ((size_t)pl.total_slots * POOL_DEEP * sizeof(int2) + 255) & ~(size_t)255;
    a.state = (float4 *)b;
    a.total_slots = pl.total_slots;
    a.n_items = n_items;
    static const int leaf_min = getenv("PTB_POOL_LEAF_MIN") ? atoi(getenv("PTB_POOL_LEAF_MIN")) : 16;
    static const int shade_min = getenv("PTB_POOL_SHADE_MIN") ? atoi(getenv("PTB_POOL_SHADE_MIN")) : 16;
    static const int node_steps = getenv("PTB_POOL_NODE_STEPS") ? atoi(getenv("PTB_POOL_NODE_STEPS")) : 2;
    a.leaf_min = leaf_min; a.shade_min = shade_min; a.node_steps = node_steps < 1 ? 1 : node_steps;
    return a;
}
bool pool_usable(const RenderParams &p) { return p.accel == PT_ACCEL_LBVH && p.sc.bvh4 != nullptr && p.sc.bvh_root >= 0 && p.sc.srec != nullptr; }

// the whole render: one memset + one launch.  Returns the number of kernels launched.
int launch_pool(const RenderParams &p, void *scratch, cudaStream_t st) {
    const long long n = (long long)p.W * p.H;
    long long items = n;
    if (p.shard_n > 1) {
        const long long g = p.shard_group, bands = (n + g - 1) / g;
        items = 0;
        for (long long b = p.shard_rank; b < bands; b += p.shard_n) items += (b + 1) * g <= n ? g : n - b * g;
    }
    if (items <= 0) return 0;
    const bool qn = p.sc.bvh4q != nullptr;
    const PoolPlan pl = p.count_work ? (qn ? pool_plan_t<true, 0, true>() : pool_plan_t<true, 0, false>()) : (qn ? pool_plan_t<false, 0, true>() : pool_plan_t<false, 0, false>());
    const PoolArgs a = pool_args(pl, scratch, (int)items);
    cudaMemsetAsync(a.cursor, 0, sizeof(int), st);
    // no more blocks than there are items to start on (a 16x16 frame does not need 1000 idle blocks)
    // (Sizing the grid so that items / slots comes out as whole rounds — fewer blocks, every slot busy to the end — was measured and is
    // slower: 2.90 vs 3.01 Grays/s on rank 0's share of an 8-way tile split, 3.81 vs 3.87 on the full 1080p frame; profiles/r02_pool_rounds.txt.)
    long long blocks = (items + POOL_WARPS * POOL_NS - 1) / (POOL_WARPS * POOL_NS);
    if (blocks > pl.blocks) blocks = pl.blocks;
    if (p.count_work) { if (qn) k_pool<true, 0, true><<<(unsigned)blocks, 32 * POOL_WARPS, pl.smem, st>>>(p, a); else k_pool<true, 0, false><<<(unsigned)blocks, 32 * POOL_WARPS, pl.smem, st>>>(p, a); }
    else { if (qn) k_pool<false, 0, true><<<(unsigned)blocks, 32 * POOL_WARPS, pl.smem, st>>>(p, a); else k_pool<false, 0, false><<<(unsigned)blocks, 32 * POOL_WARPS, pl.smem, st>>>(p, a); }
    return 1;
}
int launch_pool_closest_hit(const RenderParams &p, const RefRay *rays, size_t n, float2 *out, void *scratch, cudaStream_t st) {
    if (n == 0) return 0;
    const bool qn = p.sc.bvh4q != nullptr;
    const PoolPlan pl = qn ? pool_plan_t<false, 1, true>() : pool_plan_t<false, 1, false>();
    PoolArgs a = pool_args(pl, scratch, (int)n);
    a.rays = rays; a.out = out;
    cudaMemsetAsync(a.cursor, 0, sizeof(int), st);
    long long blocks = ((long long)n + POOL_WARPS * POOL_NS - 1) / (POOL_WARPS * POOL_NS);
    if (blocks > pl.blocks) blocks = pl.blocks;
    if (qn) k_pool<false, 1, true><<<(unsigned)blocks, 32 * POOL_WARPS, pl.smem, st>>>(p, a);
    else k_pool<false, 1, false><<<(unsigned)blocks, 32 * POOL_WARPS, pl.smem, st>>>(p, a);
    return 1;
}
