// pt_sched.cuh — phase-scheduled megakernel for the LBVH (included inside namespace PT_NS by pt_kernels.cuh).
//
// Problem (profiles/r01_lbvh_v1.txt): with one whole closest-hit query per loop iteration the warp
// waits for its slowest lane's traversal and splits between box and triangle code — 8.8 of 32 lanes
// active per instruction on the 104 k-triangle scene.
//
// Here the unit of convergence is ONE traversal step, not one ray.  Every lane is in one of four
// states and each loop iteration the warp votes and runs the phase most lanes are waiting for,
// the others predicated off:
//     NODE   the lane's next link is an internal node  -> fetch 64 B, two slab tests, descend / pop
//     LEAF   the lane's next link is a leaf            -> test its <= 8 triangles, pop
//     SHADE  traversal finished (or no ray yet)        -> shade the hit / end the path / next sample
//     DONE   all samples of the pixel rendered
// Lanes that reach a leaf or finish early simply wait until enough lanes want the same phase, so the
// expensive triangle and shading code runs with many lanes active instead of one or two.
// The result is identical to the simple loop: same closest hit (t, index) rule, same RNG order per pixel.
enum { ST_NODE = 0, ST_LEAF = 1, ST_SHADE = 2, ST_DONE = 3 };

#ifndef PT_SCHED_SMEM_STACK
#define PT_SCHED_SMEM_STACK 16
#endif
#ifndef PT_SCHED_NODE_STEPS
#define PT_SCHED_NODE_STEPS 4      // node steps per NODE phase before voting again
#endif

template <bool COUNT, int MINB>
__global__ void __launch_bounds__(PT_BLOCK, MINB) k_render_sched(const RenderParams p) {
    const SceneDev &sc = p.sc;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int y = blockIdx.y * 8 + (warp >> 1) * 4 + (lane >> 3);
    Work wk = {0, 0, 0, 0, 0, 0, 0};
    const int id = y * p.W + x;
    bool mine = x < p.W && y < p.H;
    if (mine && p.shard_n > 1) mine = ((id / p.shard_group) % p.shard_n) == p.shard_rank;
    uint32_t rng = mine ? (uint32_t)p.rnds[id] : 1u;
    V3 acc = v3(0.0f, 0.0f, 0.0f);
    if (mine && (p.sample0 != 0 || p.sum_mode)) acc = v3(p.colors[id]);
    const V3 eye = v3(p.cam.eye.x, p.cam.eye.y, p.cam.eye.z);
    const int s_end = p.sample0 + p.spp;
    int s = mine ? p.sample0 : s_end;
    const int camX = (int)p.cam.XM;
    const int cam_x = camX == p.W ? x : id % max(camX, 1), cam_y = camX == p.W ? y : id / max(camX, 1);
    int depth = 0;
    bool alive = false;                       // a path is in flight (its ray has just been traced when st == ST_SHADE)
    V3 rayP = eye, rayD = v3(0.0f, 0.0f, 1.0f);
    Path ps;
    path_reset(ps);
    // traversal state
    int st = mine ? ST_SHADE : ST_DONE;
    // traversal stack: the first PT_SCHED_SMEM_STACK entries live in shared memory (lane-interleaved, conflict
    // free), deeper entries (rare) in local memory — a 64-entry local array per thread is 196 KB per SM of
    // L1 footprint and evicts the BVH nodes (profiles/r01_lbvh_v2.txt: 338 M local sectors)
    // Each entry = (link, entry distance of that child's box): a popped entry whose box now starts beyond the
    // best hit is dropped without fetching the node.
    __shared__ int2 s_stack[PT_SCHED_SMEM_STACK * PT_BLOCK];
    int2 *my_stack = s_stack + threadIdx.x;
    int2 deep_stack[PT_LBVH_STACK - PT_SCHED_SMEM_STACK];
    int link = 0, sp = 0;
#define PT_PUSH(v, d) do { const int2 e_ = make_int2((v), __float_as_int(d)); if (sp < PT_SCHED_SMEM_STACK) my_stack[sp * PT_BLOCK] = e_; \
                           else if (sp < PT_LBVH_STACK) deep_stack[sp - PT_SCHED_SMEM_STACK] = e_; ++sp; } while (0)
    // pop until an entry survives the distance test; returns false when the stack runs empty
#define PT_POP_LIVE(ok) do { ok = false; while (sp > 0) { --sp; const int2 e_ = sp < PT_SCHED_SMEM_STACK ? my_stack[sp * PT_BLOCK] : deep_stack[sp - PT_SCHED_SMEM_STACK]; \
                             if (__int_as_float(e_.y) <= best.t) { link = e_.x; ok = true; break; } } } while (0)
    Closest best; best.t = PT_NO_HIT; best.tri = 0x7fffffff;
    float ix = 0.0f, iy = 0.0f, iz = 0.0f, ox = 0.0f, oy = 0.0f, oz = 0.0f;   // 1/D and P/D

    for (;;) {
        const unsigned bN = __ballot_sync(0xffffffffu, st == ST_NODE);
        const unsigned bL = __ballot_sync(0xffffffffu, st == ST_LEAF);
        const unsigned bS = __ballot_sync(0xffffffffu, st == ST_SHADE);
        if ((bN | bL | bS) == 0u) break;
        const int cN = __popc(bN), cL = __popc(bL), cS = __popc(bS);

        if (cN >= cL && cN >= cS) {
            // ---------------------------------------------------------------- NODE phase
#pragma unroll 1
            for (int k = 0; k < PT_SCHED_NODE_STEPS; ++k) {
                if (st == ST_NODE) {
                    const float4 *nd = sc.bvh + 4 * (size_t)link;
                    const float4 n0 = ldg4(nd), n1 = ldg4(nd + 1), n2 = ldg4(nd + 2), n3 = ldg4(nd + 3);
                    if (COUNT) wk.box += 2;
                    // slab distances as lo*inv - P*inv: one FMA per plane (boxes are inflated at build time, see lbvh.cu)
                    float a, b, l0, l1, r0, r1;
                    a = fmaf(n0.x, ix, -ox); b = fmaf(n0.y, ix, -ox); l0 = fminf(a, b); l1 = fmaxf(a, b);
                    a = fmaf(n0.z, iy, -oy); b = fmaf(n0.w, iy, -oy); l0 = fmaxf(l0, fminf(a, b)); l1 = fminf(l1, fmaxf(a, b));
                    a = fmaf(n2.x, iz, -oz); b = fmaf(n2.y, iz, -oz); l0 = fmaxf(l0, fminf(a, b)); l1 = fminf(l1, fmaxf(a, b));
                    a = fmaf(n1.x, ix, -ox); b = fmaf(n1.y, ix, -ox); r0 = fminf(a, b); r1 = fmaxf(a, b);
                    a = fmaf(n1.z, iy, -oy); b = fmaf(n1.w, iy, -oy); r0 = fmaxf(r0, fminf(a, b)); r1 = fminf(r1, fmaxf(a, b));
                    a = fmaf(n2.z, iz, -oz); b = fmaf(n2.w, iz, -oz); r0 = fmaxf(r0, fminf(a, b)); r1 = fminf(r1, fmaxf(a, b));
                    // widen the interval by 4e-6 relative on the far end and towards zero on the near end
                    const bool hl = (l0 <= l1 * 1.000004f) & (l1 >= 0.0f) & (l0 * 0.999996f <= best.t);
                    const bool hr = (r0 <= r1 * 1.000004f) & (r1 >= 0.0f) & (r0 * 0.999996f <= best.t);
                    const int ll = __float_as_int(n3.x), lr = __float_as_int(n3.y);
                    if (hl & hr) {
                        const bool left_first = l0 <= r0;
                        PT_PUSH(left_first ? lr : ll, (left_first ? r0 : l0) * 0.999996f);
                        link = left_first ? ll : lr;
                    } else if (hl | hr) {
                        link = hl ? ll : lr;
                    } else {
                        bool ok;
                        PT_POP_LIVE(ok);
                        if (!ok) st = ST_SHADE;
                    }
                    if (st == ST_NODE && link < 0) st = ST_LEAF;
                }
            }
        } else if (cL >= cS) {
            // ---------------------------------------------------------------- LEAF phase
            if (st == ST_LEAF) {
                const int enc = ~link;
                const int first = enc >> 3, count = (enc & 7) + 1;
                for (int k = 0; k < count; ++k) {
                    const int pos = first + k;
                    float t;
                    const bool acc_ = tri_accept<COUNT>(ldg4(sc.sq0 + pos), ldg4(sc.sq1 + pos), ldg4(sc.sq2 + pos), rayP, rayD, t, wk);
                    if (acc_ & (t > 0.0f) & (t <= best.t)) {
                        const int tid = __ldg(sc.sorted_id + pos);
                        if (t < best.t || tid < best.tri) { best.t = t; best.tri = tid; }
                    }
                }
                bool ok;
                PT_POP_LIVE(ok);
                st = !ok ? ST_SHADE : (link >= 0 ? ST_NODE : ST_LEAF);
            }
        } else {
            // ---------------------------------------------------------------- SHADE phase
            if (st == ST_SHADE) {
                if (alive) {                                         // the ray of this path has been traced
                    wk.rays++;
                    if (best.t < PT_NO_HIT) {
                        V3 hitP = rayP + rayD * best.t;              // prog.cl:107
                        V3 hitN = tri_normal(ldg4(sc.q0 + best.tri), ldg4(sc.q1 + best.tri), ldg4(sc.q2 + best.tri));
                        Mat m = load_material(sc.mats, (int)__ldg(sc.mati + best.tri));
                        shade_hit(m, hitP, hitN, rayP, rayD, rng, ps, eye, p.iterations);
                        if (++depth >= p.iterations) alive = false;
                    } else {
                        alive = false;                               // miss: break, prog.cl:367-376
                    }
                    if (!alive) {                                    // sample finished, prog.cl:379
                        if (p.sum_mode) acc = acc + ps.color;
                        else acc = pt_mean_step(acc, ps.color, s);
                        ++s;
                        wk.samples++;
                    }
                }
                while (!alive && s < s_end) {                        // next sample of this pixel
                    if (p.from_rays) {
                        const float4 *r = reinterpret_cast<const float4 *>(p.rays + id);
                        rayP = v3(r[0]); rayD = v3(r[1]);
                    } else {
                        float r1 = pt_rand(rng), r2 = pt_rand(rng);
                        camera_ray(p, cam_x, cam_y, r1, r2, rayP, rayD);
                    }
                    path_reset(ps);
                    depth = 0;
                    alive = true;
                    if (p.iterations <= 0) {                         // the bounce loop body never runs
                        alive = false;
                        if (p.sum_mode) acc = acc + ps.color; else acc = pt_mean_step(acc, ps.color, s);
                        ++s;
                        wk.samples++;
                    }
                }
                if (alive) {                                         // start the closest-hit query of the (new) ray
                    ix = safe_rcp(rayD.x); iy = safe_rcp(rayD.y); iz = safe_rcp(rayD.z);
                    ox = rayP.x * ix; oy = rayP.y * iy; oz = rayP.z * iz;
                    best.t = PT_NO_HIT; best.tri = 0x7fffffff;
                    lbvh_big_scan<COUNT>(sc, rayP, rayD, best, wk);
                    sp = 0;
                    link = sc.bvh_root;
                    st = link >= 0 ? ST_NODE : ST_LEAF;
                } else {
                    st = ST_DONE;
                }
            }
        }
    }
    if (mine) {
        p.rnds[id] = (int)rng;
        p.colors[id] = make_float4(acc.x, acc.y, acc.z, 0.0f);
        if (p.tex) p.tex[id] = reinhard_tone_map(acc);
    }
    if (p.counters) {
        unsigned v[7] = {wk.rays, wk.box, wk.tri, wk.tri1, wk.tri2, wk.tri3, wk.samples};
        const int idx[7] = {PT_CNT_RAYS, PT_CNT_BOX, PT_CNT_TRI, PT_CNT_TRI1, PT_CNT_TRI2, PT_CNT_TRI3, PT_CNT_PIXEL_SAMPLES};
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            if (!COUNT && k >= 1 && k <= 5) continue;
            unsigned long long tot = v[k];
            for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
            if (lane == 0 && tot) atomicAdd(p.counters + idx[k], tot);
        }
    }
}
