// emu_update.cpp — CPU emulation of the CUDA `update` kernels for machines without a GPU.
// TEST INFRASTRUCTURE: compiles the product's own per-ray code (gpu_ray_tracing_b200/csrc/rt_core.cuh)
// for the host and drives it with the same broad-phase / narrow-phase control flow as
// rt_kernels.cu, so that `-m "not gpu"` tests can (a) compare the kernels' arithmetic bit-for-bit
// with the oracle and (b) audit the broad phase: every (ray, sphere) pair the broad phase rejects
// is also put through the exact quadratic, and a pair it wrongly rejected counts as a violation.
// Built with -ffp-contract=off; fmaf() is the IEEE fused operation, as FFMA is on the GPU.
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <utility>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../../gpu_ray_tracing_b200/csrc/rt_core.cuh"
#include "../../gpu_ray_tracing_b200/csrc/rt_tree.h"

using rtc::f3;

extern "C" __attribute__((visibility("default")))
void emu_update(const SceneCamera* cam_p, const GpuSphere* spheres, uint32_t count, uint32_t width, uint32_t height,
                uint32_t row_begin, uint32_t row_count, float* accum,
                uint64_t* stats /* rays, exact, violations, primary rays, untraced, (tile, record) pairs kept, ray candidates a tile dropped */, int variant) {
    (void)height;
    const SceneCamera& cam = *cam_p;
    std::vector<float> gk(count);
    for (uint32_t i = 0; i < count; ++i) gk[i] = rtc::broad_record_k(spheres[i]);
    // common-origin primary pass: same applicability rule as rtk::primary_pass_applies (variant 0 / 4)
    const double rho = rtc::primary_rho(cam);
    const bool primary_ok = (variant == 0 || variant == 4) && count > 0 && rho >= 0.0 && rtc::f2u(cam.max_depth) > 0u;
    // the tree of bounding spheres (rt_tree.h): the primary pass walks it per pixel tile with this frame's records
    // (rtk::prep_primary_tree_kernel), the bounce rays per ray (rtk::trace_tree) when its image fits shared memory
    rtt::Tree tree;
    const bool primary = primary_ok && rtt::build(spheres, count, tree);
    const bool clustered = primary && count < 32768u && tree.bounce_bytes() <= (size_t)2 * (2048 + 8) * 16 &&
                           tree.n_nodes < 65536u && 16u + 4u + 7u * tree.depth <= 64u;
    std::vector<float> ptree(primary ? (size_t)tree.n_nodes * rtt::kStride * 4 : 0);   // per child slot
    std::vector<float> prim(primary ? (size_t)count * 4 : 0);                          // per sphere (= its leaf slot's record)
    if (primary) {
        const float m_cam = rtc::primary_m_cam(cam, rho);
        for (uint32_t node = 0; node < tree.n_nodes; ++node)
            for (int k = 0; k < rtt::kStride; ++k) {
                float* out = &ptree[((size_t)node * rtt::kStride + k) * 4];
                rtc::primary_record_pad(out);
                if (k >= rtt::kWide) continue;
                const uint32_t lk = tree.link[(size_t)node * rtt::kWide + k];
                if (lk == rtt::kNone) continue;
                const float* b = &tree.rec[((size_t)node * rtt::kStride + k) * 4];
                if (node >= tree.first_leaf) {
                    rtc::primary_record(spheres[lk], rtc::mk(cam.center), rho, out);
                    memcpy(&prim[(size_t)lk * 4], out, 16);
                } else if (b[3] < -1.0e38f) {
                    out[0] = 0.0f; out[1] = 0.0f; out[2] = 0.0f; out[3] = 1.0f;
                } else {
                    rtc::primary_group_record(b[0], b[1], b[2], tree.child_r[(size_t)node * rtt::kWide + k], m_cam, rtc::mk(cam.center), rho, out);
                }
            }
    }
    uint64_t cluster_rays = 0, cluster_tests = 0, member_tests = 0, retraced = 0;
    uint64_t prim_rays = 0, untraced = 0, tile_pairs = 0, tile_dropped = 0, tile_steps = 0;
    const uint32_t seed_bits_early = rtc::f2u(cam.random_seed * 4294967296.0f);
    // tile level of the primary pass (rtk::prim_tile_bound / prim_cull): per 32 x 8 pixel tile of the window, the
    // records that survive the tile test.  tile_cand[tile * count + s] != 0 <=> record s is scanned for that tile.
    const uint32_t tiles_x = (width + 31u) / 32u, tiles_y = (row_count + 7u) / 8u;
    std::vector<uint8_t> tile_cand(primary ? (size_t)tiles_x * tiles_y * count : 0);
    // ... and the same records as the warp scans them: nearest first by rtc::primary_near (rtk::prim_sort_list)
    std::vector<std::vector<std::pair<float, uint32_t>>> tile_list(primary ? (size_t)tiles_x * tiles_y : 0);
    if (primary) {
        const uint32_t spp_e = rtc::f2u(cam.samples_per_pixel);
        const bool reset_e = cam.camera_has_moved > 0.5f;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : tile_pairs, tile_steps)
#endif
        for (int64_t tile = 0; tile < (int64_t)tiles_x * tiles_y; ++tile) {
            const uint32_t x0 = (uint32_t)(tile % tiles_x) * 32u, r0 = (uint32_t)(tile / tiles_x) * 8u;
            float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
            std::vector<float> us;
            for (uint32_t row = r0; row < r0 + 8u && row < row_count; ++row)
                for (uint32_t x = x0; x < x0 + 32u && x < width; ++x) {
                    const float* texel = accum + ((size_t)row * width + x) * 4;
                    const uint32_t n = reset_e ? 0u : rtc::f2u(texel[3]);
                    if (!(n < spp_e)) continue;
                    const uint32_t seed = 1u + n + seed_bits_early;
                    const rtc::Ray r = rtc::get_ray(cam, (int)x, (int)(row_begin + row), seed, seed_bits_early);
                    float u[3];
                    if (!rtc::primary_arm(r.d, rtc::dot(r.d, r.d), &u[0], &u[1], &u[2])) continue;
                    for (int k = 0; k < 3; ++k) {
                        lo[k] = fminf(lo[k], u[k]);
                        hi[k] = fmaxf(hi[k], u[k]);
                        us.push_back(u[k]);
                    }
                }
            if (us.empty()) continue;
            const f3 m = rtc::primary_tile_axis(rtc::mk(lo[0], lo[1], lo[2]), rtc::mk(hi[0], hi[1], hi[2]));
            float e2 = 0.0f;
            for (size_t k = 0; k < us.size(); k += 3) e2 = fmaxf(e2, rtc::primary_tile_e2(us[k], us[k + 1], us[k + 2], m));
            const float chord = rtc::primary_tile_chord(e2);
            std::vector<uint32_t> work;   // the tile's walk: the entry list, then down the tree
            for (uint32_t e = 0; e < tree.n_entry; ++e) {
                const uint32_t slot = tree.entry_slot[e];
                ++tile_steps;
                if (slot != rtt::kNone) {
                    const float* rec = &ptree[(size_t)slot * 4];
                    if (!rtc::primary_tile_candidate(rec[0], rec[1], rec[2], rec[3], m, chord)) continue;
                }
                work.push_back(tree.entry_node[e]);
            }
            while (!work.empty()) {
                const uint32_t node = work.back();
                work.pop_back();
                for (int k = 0; k < rtt::kWide; ++k) {
                    const uint32_t lk = tree.link[(size_t)node * rtt::kWide + k];
                    if (lk == rtt::kNone) continue;
                    const float* rec = &ptree[((size_t)node * rtt::kStride + k) * 4];
                    ++tile_steps;
                    if (!rtc::primary_tile_candidate(rec[0], rec[1], rec[2], rec[3], m, chord)) continue;
                    if (node >= tree.first_leaf) {
                        tile_cand[(size_t)tile * count + lk] = 1;
                        tile_list[(size_t)tile].push_back({rtc::primary_near(rec[0], rec[1], rec[2], rec[3]), lk});
                        ++tile_pairs;
                    } else {
                        work.push_back(lk);
                    }
                }
            }
            std::stable_sort(tile_list[(size_t)tile].begin(), tile_list[(size_t)tile].end(),
                             [](const std::pair<float, uint32_t>& x, const std::pair<float, uint32_t>& y) { return x.first < y.first; });
        }
    }
    const uint32_t seed_bits = rtc::f2u(cam.random_seed * 4294967296.0f);
    const uint32_t spp = rtc::f2u(cam.samples_per_pixel);
    const uint32_t max_depth = rtc::f2u(cam.max_depth);
    const bool reset = cam.camera_has_moved > 0.5f;
    uint64_t rays = 0, exact = 0, violations = 0, beaten = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : rays, exact, violations, beaten, prim_rays, untraced, tile_dropped, cluster_rays, cluster_tests, member_tests, retraced)
#endif
    for (int64_t row = 0; row < (int64_t)row_count; ++row) {
        const int y_abs = (int)(row_begin + row);
        std::vector<uint32_t> stamp(primary ? count : 0, 0u);   // [s] == ray number: s went through (or was culled before) the exact test
        uint32_t ray_no = 0;
        for (uint32_t x = 0; x < width; ++x) {
            float* texel = accum + ((size_t)row * width + x) * 4;
            f3 mean = rtc::mk(texel[0], texel[1], texel[2]);
            uint32_t n = rtc::f2u(texel[3]);
            if (reset) { mean = rtc::mk(0, 0, 0); n = 0; }
            if (n < spp) {
                const uint32_t seed = 1u + n + seed_bits;
                rtc::Ray r = rtc::get_ray(cam, (int)x, y_abs, seed, seed_bits);
                f3 factor = rtc::mk(1.0f, 1.0f, 1.0f);
                bool black = false;
                for (uint32_t i = 0; i < max_depth; ++i) {
                    ++rays;
                    const float a = rtc::dot(r.d, r.d);
                    const rtc::BroadRay b = rtc::broad_arm(r.o, r.d, a);
                    float best_t = 3.4e35f;
                    int best_i = -1;
                    float pux, puy, puz, inv_len = 0.0f;
                    const bool use_primary = primary && i == 0;
                    if (use_primary && rtc::primary_arm(r.d, a, &pux, &puy, &puz, &inv_len)) {
                        ++prim_rays;
                        ++ray_no;
                        const size_t tile = (size_t)(row / 8) * tiles_x + x / 32u;
                        const uint8_t* tc = &tile_cand[tile * count];
                        // the kernel culls with the smallest 1/|d| and the farthest hit of a strip of 8 rays; this ray's own
                        // values cull at least as much, so the audit below covers every cull the kernel can make
                        const float near_scale = rtc::primary_near_scale(inv_len);
                        bool nan_root = false;
                        for (const std::pair<float, uint32_t>& item : tile_list[tile]) {   // nearest first
                            const uint32_t s = item.second;
                            const GpuSphere& sp = spheres[s];
                            const float* rec = &prim[(size_t)s * 4];
                            if (!rtc::primary_candidate(rtc::primary_h(rec[0], rec[1], rec[2], rec[3], pux, puy, puz))) continue;
                            stamp[s] = ray_no;
                            if (item.first * near_scale > best_t) {   // already beaten (rtk::prim_scan_list)
                                ++beaten;
                                float root;
                                if (rtc::sphere_root(r.o, r.d, a, rtc::mk(sp.position), sp.radius, 0.001f, 3.4e35f, &root) &&
                                    !(root > best_t))
                                    ++violations;   // ... but the reference's root is not beyond the hit held (NaN included)
                                continue;
                            }
                            ++exact;
                            float root;
                            if (rtc::sphere_root(r.o, r.d, a, rtc::mk(sp.position), sp.radius, 0.001f, 3.4e35f, &root) &&
                                root < item.first * near_scale)
                                ++violations;   // the bound primary_near rests on fails for a root the reference accepts
                                                // (a NaN root belongs to an always-record, whose bound is 0: never culled)
                            if (!rtc::nearest_update(r.o, r.d, a, rtc::mk(sp.position), sp.radius, 0.001f, 3.4e35f, (int)s, &best_t, &best_i))
                                nan_root = true;
                        }
                        for (uint32_t s = 0; s < count; ++s) {
                            if (stamp[s] == ray_no) continue;
                            const GpuSphere& sp = spheres[s];
                            const float* rec = &prim[(size_t)s * 4];
                            float root;
                            const bool ray_cand = rtc::primary_candidate(rtc::primary_h(rec[0], rec[1], rec[2], rec[3], pux, puy, puz));
                            if (ray_cand && !tc[s]) ++tile_dropped;   // expected 0: the tile level keeps what any ray test keeps
                            if (rtc::sphere_root(r.o, r.d, a, rtc::mk(sp.position), sp.radius, 0.001f, 3.4e35f, &root))
                                ++violations;   // rejected a sphere the reference hits inside the original (tmin, tmax)
                        }
                        if (nan_root) {   // the kernel hands such a ray to its sequential fallback
                            best_t = 3.4e35f;
                            best_i = -1;
                            for (uint32_t s = 0; s < count; ++s) {
                                float root;
                                if (rtc::sphere_root(r.o, r.d, a, rtc::mk(spheres[s].position), spheres[s].radius, 0.001f, best_t, &root)) {
                                    best_t = root;
                                    best_i = (int)s;
                                }
                            }
                        }
                    } else if (clustered && i >= 1) {
                        ++cluster_rays;
                        const float m_ray = rtc::cluster_m(rtc::dot(r.o, r.o));
                        std::vector<uint8_t> tested(count, 0);
                        bool nan_root = false;
                        std::vector<uint32_t> stack;
                        {   // the coherent scan of the entry list
                            const float t_entry = rtc::cluster_thr(b.thr, m_ray, tree.entry_r);
                            for (uint32_t e = 0; e < tree.n_entry; ++e) {
                                const float* rec = &tree.entry_rec[(size_t)e * 4];
                                ++cluster_tests;
                                if (!(rtc::broad_g(rec[0], rec[1], rec[2], rec[3], b.ux, b.uy, b.uz, b.h0, b.mox, b.moy, b.moz) < t_entry))
                                    stack.push_back(tree.entry_node[e]);
                            }
                        }
                        while (!stack.empty()) {
                            const uint32_t node = stack.back();
                            stack.pop_back();
                            const bool leaf = node >= tree.first_leaf;
                            const float t = leaf ? b.thr : rtc::cluster_thr(b.thr, m_ray, tree.node_r[node]);
                            for (int k = 0; k < rtt::kWide; ++k) {
                                const float* rec = &tree.rec[((size_t)node * rtt::kStride + k) * 4];
                                const uint32_t link = tree.link[(size_t)node * rtt::kWide + k];
                                ++(leaf ? member_tests : cluster_tests);
                                if (rtc::broad_g(rec[0], rec[1], rec[2], rec[3], b.ux, b.uy, b.uz, b.h0, b.mox, b.moy, b.moz) < t) continue;
                                if (link == rtt::kNone) continue;
                                if (!leaf) {
                                    stack.push_back(link);
                                    continue;
                                }
                                const GpuSphere& sp = spheres[link];
                                ++exact;
                                tested[link] = 1;
                                if (!rtc::nearest_update_from_inside(r.o, r.d, a, rtc::mk(sp.position), sp.radius, 0.001f, 3.4e35f, (int)link,
                                                         &best_t, &best_i))
                                    nan_root = true;
                            }
                        }
                        for (uint32_t s = 0; s < count; ++s) {
                            float root;
                            if (!tested[s] && rtc::sphere_root(r.o, r.d, a, rtc::mk(spheres[s].position), spheres[s].radius, 0.001f,
                                                               3.4e35f, &root))
                                ++violations;   // the tree skipped a sphere the reference hits
                        }
                        if (nan_root) {   // order matters once a root is NaN: the reference loop itself
                            ++retraced;
                            best_t = 3.4e35f;
                            best_i = -1;
                            for (uint32_t s = 0; s < count; ++s) {
                                float root;
                                if (rtc::sphere_root(r.o, r.d, a, rtc::mk(spheres[s].position), spheres[s].radius, 0.001f, best_t, &root)) {
                                    best_t = root;
                                    best_i = (int)s;
                                }
                            }
                        }
                    } else {
                    if (use_primary) ++untraced;
                    for (uint32_t s = 0; s < count; ++s) {
                        const GpuSphere& sp = spheres[s];
                        const float g = rtc::broad_g(sp.position.x, sp.position.y, sp.position.z, gk[s], b.ux, b.uy, b.uz,
                                                     b.h0, b.mox, b.moy, b.moz);
                        float root;
                        if (!(g < b.thr)) {
                            ++exact;
                            if (rtc::sphere_root(r.o, r.d, a, rtc::mk(sp.position), sp.radius, 0.001f, best_t, &root)) {
                                best_t = root;
                                best_i = (int)s;
                            }
                        } else if (rtc::sphere_root(r.o, r.d, a, rtc::mk(sp.position), sp.radius, -INFINITY, INFINITY, &root)) {
                            ++violations;   // broad phase rejected a pair whose exact discriminant is >= 0 (or NaN)
                        }
                    }
                    }
                    if (best_i < 0) break;
                    if (!rtc::scatter(&r, &factor, best_t, spheres[best_i], rtc::hash(seed + 1u + i * 1000u))) {
                        black = true;
                        break;
                    }
                }
                f3 color = black ? rtc::mk(0, 0, 0) : rtc::sky_times(factor, r.d);
                mean = rtc::add(mean, rtc::divs(rtc::sub(color, mean), (float)(n + 1u)));
                n += 1u;
            }
            texel[0] = mean.x; texel[1] = mean.y; texel[2] = mean.z; texel[3] = (float)n;
        }
    }
    if (stats) {
        stats[12] = beaten;
        stats[0] = rays; stats[1] = exact; stats[2] = violations; stats[3] = prim_rays; stats[4] = untraced;
        stats[5] = tile_pairs; stats[6] = tile_dropped;
        stats[11] = tile_steps;
        stats[7] = cluster_rays; stats[8] = cluster_tests; stats[9] = member_tests; stats[10] = retraced;
    }
}

// Direct access to the primary-pass record functions for invariant tests.
extern "C" __attribute__((visibility("default")))
double emu_primary_rho(const SceneCamera* cam) { return rtc::primary_rho(*cam); }

extern "C" __attribute__((visibility("default")))
void emu_primary_record(const SceneCamera* cam, const GpuSphere* sphere, float* out4) {
    rtc::primary_record(*sphere, rtc::mk(cam->center), rtc::primary_rho(*cam), out4);
}
