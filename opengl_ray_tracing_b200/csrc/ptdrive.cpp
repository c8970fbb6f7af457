// ptdrive — the reference's main() / display() re-hosted as a headless program over the two C ABIs.
//
// What main() of ImportanceSampling_LowDiscrepancySequence/main.cpp ("M:") does, in its order, with every GL call replaced
// by the libpt.so entry point INTEGRATION.md names for it:
//   M:752-777  materials + readObj of the model and of the ground quad      -> pts_read_obj / pts_add_icosphere / pts_add_quad
//   M:779-790  dummy node + buildBVHwithSAH(…, 8)                            -> pts_build_bvh
//   M:792-821  encode loops                                                  -> pts_triangles_encoded / pts_nodes_encoded
//   M:843-853  HDRLoader::load + calculateHdrCache + two glTexImage2D        -> pts_hdr_load / pts_hdr_cache, pt_upload_scene
//   M:659-709  display(): camera uniforms, frameCounter++, pass1/pass2/pass3 -> pt_set_camera, pt_render_spp, pt_readback_display
// and, because there is no window, the image pass3 would show is written as a PNG (and the fp32 radiance as a PFM).
//
// --gpus N renders on N GPUs (one context and one host thread per GPU, the scene replicated): the image is split in diagonal
// 8x4 tiles (--split tile: bit-identical to one GPU) or the frames are dealt round-robin (--split sample), and pt_combine
// merges the N framebuffers into rank 0's image over NVLink.
//
//   ptdrive --variant 2 --icosphere 71 --size 1920x1080 --spp 256 --out image.png
//   ptdrive --variant 2 --icosphere 224 --size 1920x1080 --spp 4096 --gpus 8 --out image.png
//   ptdrive --variant 0 --obj model.obj --hdr sky.hdr --spp 64 --camera 10,25,2.5 --out image.png --pfm radiance.pfm
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/pt.h"
#include "../../include/ptscene.h"

namespace {

struct Options {
    int variant = 2;
    std::string obj, hdr, out = "ptdrive.png", pfm;
    int icosphere = 22;                 // 20 n^2 triangles when no --obj is given
    int width = 512, height = 512;      // M:82-83
    int hdr_w = 1024, hdr_h = 512;      // procedural sky when no --hdr is given
    unsigned spp = 64, per_call = 64;
    float up = 0.0f, rot = 0.0f, r = 4.0f;   // M:147-149
    int sah = 1, threads = 0, mode = PT_MODE_FAST, pipeline = PT_PIPELINE_PERSISTENT, device = 0, max_bounce = 0;
    bool basic = false;                 // --basic: BasicRayTracingWithC++/main.cpp instead (its own scene literal, 256x256)
    unsigned seed = 1;
    bool size_given = false;
    int gpus = 1;                       // --gpus N: contexts on devices device .. device + N - 1
    bool split_tiles = true;            // --split tile | sample
};

[[noreturn]] void die(const std::string& msg) {
    std::fprintf(stderr, "ptdrive: %s\n", msg.c_str());
    std::exit(1);
}

void usage() {
    std::puts("ptdrive [--variant 0|1|2] [--obj FILE | --icosphere N] [--hdr FILE | --synth-hdr WxH] [--size WxH] [--spp N]\n"
              "        [--spp-per-call K] [--camera UP,ROT,R] [--median] [--threads T] [--mode M] [--pipeline P] [--device D]\n"
              "        [--max-bounce B] [--gpus N [--split tile|sample]] [--out image.png] [--pfm radiance.pfm]\n"
              "ptdrive --basic [--size WxH] [--spp N] [--seed S] [--device D] [--out image.png]      (BasicRayTracingWithC++/main.cpp)");
}

bool parse_pair(const char* s, char sep, int& a, int& b) {
    const char* p = std::strchr(s, sep);
    if (!p) return false;
    a = std::atoi(s); b = std::atoi(p + 1);
    return a > 0 && b > 0;
}

Options parse(int argc, char** argv) {
    Options o;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&]() -> const char* { if (i + 1 >= argc) die("missing value after " + a); return argv[++i]; };
        if (a == "--variant") o.variant = std::atoi(next());
        else if (a == "--obj") o.obj = next();
        else if (a == "--icosphere") o.icosphere = std::atoi(next());
        else if (a == "--hdr") o.hdr = next();
        else if (a == "--synth-hdr") { if (!parse_pair(next(), 'x', o.hdr_w, o.hdr_h)) die("--synth-hdr wants WxH"); }
        else if (a == "--size") { if (!parse_pair(next(), 'x', o.width, o.height)) die("--size wants WxH"); o.size_given = true; }
        else if (a == "--basic") o.basic = true;
        else if (a == "--seed") o.seed = (unsigned)std::strtoul(next(), nullptr, 10);
        else if (a == "--spp") o.spp = (unsigned)std::atoi(next());
        else if (a == "--spp-per-call") o.per_call = (unsigned)std::atoi(next());
        else if (a == "--camera") { if (std::sscanf(next(), "%f,%f,%f", &o.up, &o.rot, &o.r) != 3) die("--camera wants UP,ROT,R"); }
        else if (a == "--median") o.sah = 0;
        else if (a == "--threads") o.threads = std::atoi(next());
        else if (a == "--mode") o.mode = std::atoi(next());
        else if (a == "--pipeline") o.pipeline = std::atoi(next());
        else if (a == "--device") o.device = std::atoi(next());
        else if (a == "--max-bounce") o.max_bounce = std::atoi(next());
        else if (a == "--gpus") o.gpus = std::atoi(next());
        else if (a == "--split") { const std::string v = next(); if (v != "tile" && v != "sample") die("--split wants tile or sample"); o.split_tiles = v == "tile"; }
        else if (a == "--out") o.out = next();
        else if (a == "--pfm") o.pfm = next();
        else if (a == "-h" || a == "--help") { usage(); std::exit(0); }
        else die("unknown option " + a);
    }
    if (o.variant < 0 || o.variant > 2) die("--variant must be 0, 1 or 2");
    if (o.spp == 0 || o.per_call == 0) die("--spp and --spp-per-call must be positive");
    if (o.gpus < 1 || o.gpus > 16) die("--gpus must be 1 .. 16");
    return o;
}

double now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

void ptcheck(int rc, const char* what) { if (rc != PT_OK) die(std::string(what) + ": " + pt_last_error()); }
void ptscheck(int rc, const char* what) { if (rc < 0) die(std::string(what) + ": " + pts_last_error()); }

// The scene each reference main() builds around its model: material literals and transforms of the three programs.
void populate(pts_scene* sc, const Options& o) {
    const float zero[3] = {0, 0, 0};
    float T[16];
    PtsMaterial m;
    pts_material_default(&m);
    auto mesh = [&](const PtsMaterial& mat, const float trans[3], float s, unsigned seed) {
        const float scale[3] = {s, s, s};
        pts_get_transform_matrix(zero, trans, scale, T);
        if (!o.obj.empty()) ptscheck(pts_read_obj(sc, o.obj.c_str(), &mat, T, 1), "readObj");
        else ptscheck(pts_add_icosphere(sc, o.icosphere, 0.15f, seed, &mat, T, 1), "icosphere");
    };
    auto quad = [&](const PtsMaterial& mat, float y, float len) {
        const float trans[3] = {0, y, 0}, scale[3] = {len, 0.01f, len};
        pts_get_transform_matrix(zero, trans, scale, T);
        ptscheck(pts_add_quad(sc, &mat, T), "quad");
    };
    if (o.variant == PT_VARIANT_LAMBERT) {                        // OpenglRayTracing/main.cpp:650-670
        m.baseColor[0] = 0; m.baseColor[1] = 1; m.baseColor[2] = 1;
        const float t[3] = {0.3f, -0.4f, 0};
        mesh(m, t, 1.5f, 1);
        pts_material_default(&m);
        m.baseColor[0] = 0.725f; m.baseColor[1] = 0.71f; m.baseColor[2] = 0.68f;
        quad(m, -1.4f, 18.83f);
        pts_material_default(&m);
        m.emissive[0] = m.emissive[1] = m.emissive[2] = 20.0f;
        const float lt[3] = {0, 1.38f, 0}, ls[3] = {0.7f, 0.01f, 0.7f};
        pts_get_transform_matrix(zero, lt, ls, T);
        ptscheck(pts_add_quad(sc, &m, T), "light quad");
    } else if (o.variant == PT_VARIANT_DISNEY_UNIFORM) {                  // DisneyBRDF/main.cpp:720-731
        m.baseColor[0] = 0.75f; m.baseColor[1] = 0.7f; m.baseColor[2] = 0.15f;
        m.roughness = 0.15f; m.metallic = 1.0f; m.specular = 0.5f; m.clearcoat = 1.0f;
        const float t[3] = {0, -0.4f, 0};
        mesh(m, t, 1.75f, 2);
    } else {                                                      // M:752-777
        m.baseColor[0] = 1; m.baseColor[1] = 0.73f; m.baseColor[2] = 0.25f;
        m.roughness = 0.5f; m.specular = 1.0f; m.metallic = 1.0f; m.clearcoat = 1.0f; m.clearcoatGloss = 0.0f;
        const float t[3] = {0, -0.5f, 0};
        mesh(m, t, 0.75f, 3);
        m.baseColor[0] = m.baseColor[1] = m.baseColor[2] = 1;     // the reference mutates `m` in place (M:766-769)
        m.roughness = 0.01f; m.metallic = 0.1f;
        quad(m, -0.5f, 13.0f);
    }
}

// BasicRayTracingWithC++/main.cpp: scene literal, SAMPLE passes, imshow -> PNG (B:300-441)
int run_basic(const Options& o) {
    const int w = o.size_given ? o.width : 256, h = o.size_given ? o.height : 256;      // B:23-24
    PtBasicShape shapes[64];
    const int n = pt_basic_default_scene(shapes, 64);
    std::vector<double> image((size_t)w * h * 3);
    double ms = 0.0;
    const double t0 = now();
    if (pt_basic_render(o.device, shapes, n, w, h, (int)o.spp, o.seed, image.data(), &ms) != PT_OK) die(std::string("pt_basic_render: ") + pt_basic_last_error());
    std::printf("finish in %.3f seconds (%.2f ms on the device), %d shapes, %dx%d, %u samples, %.1f Mpaths/s\n", now() - t0, ms, n, w, h, o.spp,
                (double)w * h * o.spp / (ms * 1e3));
    std::vector<unsigned char> rgb((size_t)w * h * 3);
    if (pt_basic_imshow(image.data(), w, h, rgb.data()) != PT_OK) die(pt_basic_last_error());
    ptscheck(pts_write_png(o.out.c_str(), rgb.data(), w, h, 3), "pts_write_png");
    std::printf("wrote %s\n", o.out.c_str());
    return 0;
}

// display() on N GPUs: one host thread and one context per GPU; rank r renders its share, pt_combine merges.
void run_multi_gpu(const Options& o, pts_scene* sc, int nTriangles, int nNodes, const float* hdr, const float* cache, int hw, int hh,
                   const float eye[3], const float cam[16]) {
    const int N = o.gpus;
    unsigned char id[PT_COMM_ID_BYTES];
    ptcheck(pt_comm_get_unique_id(id, sizeof id), "pt_comm_get_unique_id");
    std::vector<unsigned char> rgba8((size_t)o.width * o.height * 4);
    std::vector<float> rgba(o.pfm.empty() ? 0 : (size_t)o.width * o.height * 4);
    std::vector<PtStats> stats((size_t)N);
    std::vector<double> wall((size_t)N, 0.0);
    std::vector<int> peer((size_t)N, 0);
    auto worker = [&](int r) {
        // a rank that fails ends the process at once: the other ranks would wait for it in the next collective for ever
        auto check = [&](int rc, const char* what) { if (rc != PT_OK) die("rank " + std::to_string(r) + ": " + what + ": " + pt_last_error()); return true; };
        PtConfig cfg = {o.variant, o.width, o.height, o.max_bounce, o.device + r, o.mode, o.pipeline};
        pt_ctx* pt = nullptr;
        if (!check(pt_create(&cfg, &pt), "pt_create")) return;
        bool ok = check(pt_upload_scene(pt, pts_triangles_encoded(sc), nTriangles, pts_nodes_encoded(sc), nNodes, hdr, cache, hw, hh), "pt_upload_scene") &&
                  check(pt_comm_init(pt, N, r, id, sizeof id), "pt_comm_init") && check(pt_set_camera(pt, eye, cam), "pt_set_camera");
        if (ok && o.split_tiles) ok = check(pt_set_tile_interleave(pt, N, r), "pt_set_tile_interleave");
        if (ok) pt_comm_info(pt, nullptr, nullptr, &peer[(size_t)r]);
        const double t0 = now();
        unsigned done = 0;
        while (ok && done < o.spp) {                                     // the display() loop, this rank's share of every call
            const unsigned n = o.spp - done < o.per_call ? o.spp - done : o.per_call;
            if (o.split_tiles) {
                ok = check(pt_render_spp(pt, done, n), "pt_render_spp");
            } else {
                const unsigned mine = n > (unsigned)r ? (n - (unsigned)r + (unsigned)N - 1) / (unsigned)N : 0;   // frames done + r, done + r + N, ...
                if (mine) ok = check(pt_render_frames(pt, done + (unsigned)r, mine, (unsigned)N, PT_ACC_SUM), "pt_render_frames");
            }
            done += n;
        }
        const int flags = (o.split_tiles ? PT_COMBINE_DISJOINT : PT_COMBINE_SUM) | PT_COMBINE_DISPLAY;
        if (ok) ok = check(pt_combine(pt, flags, o.spp, r == 0 && !rgba.empty() ? rgba.data() : nullptr, r == 0 ? rgba8.data() : nullptr), "pt_combine");
        if (ok) ok = check(pt_synchronize(pt), "pt_synchronize");
        wall[(size_t)r] = now() - t0;
        if (ok) check(pt_get_stats(pt, &stats[(size_t)r]), "pt_get_stats");
        pt_destroy(pt);
    };
    std::vector<std::thread> th;
    for (int r = 0; r < N; r++) th.emplace_back(worker, r);
    for (auto& t : th) t.join();
    double ms = 0.0, w = 0.0;
    unsigned long long rays = 0;
    for (int r = 0; r < N; r++) { ms = stats[(size_t)r].render_ms > ms ? stats[(size_t)r].render_ms : ms; w = wall[(size_t)r] > w ? wall[(size_t)r] : w; rays += stats[(size_t)r].rays; }
    std::printf("rendered %u spp at %dx%d on %d GPUs (%s split, combine over %s): %.2f ms on the slowest device (%.2f ms wall), %.1f Mrays/s\n", o.spp, o.width,
                o.height, N, o.split_tiles ? "tile" : "sample", peer[0] ? "NVLink peer memory" : "ncclReduce", ms, w * 1e3, (double)rays / (ms * 1e3));
    ptscheck(pts_write_png(o.out.c_str(), rgba8.data(), o.width, o.height, 4), "pts_write_png");
    std::printf("wrote %s\n", o.out.c_str());
    if (!o.pfm.empty()) {
        ptscheck(pts_write_pfm(o.pfm.c_str(), rgba.data(), o.width, o.height), "pts_write_pfm");
        std::printf("wrote %s\n", o.pfm.c_str());
    }
}

}  // namespace

int main(int argc, char** argv) {
    const Options o = parse(argc, argv);
    std::printf("%s\n", pt_build_info());
    if (o.basic) return run_basic(o);

    // ---- scene, BVH, encoded arrays (M:752-821) ----
    double t0 = now();
    pts_scene* sc = pts_scene_create();
    populate(sc, o);
    const int threads = o.threads > 0 ? o.threads : 16;
    const int nNodes = pts_build_bvh(sc, 8, o.sah, threads);
    ptscheck(nNodes, "buildBVH");
    const int nTriangles = pts_triangle_count(sc);
    std::printf("scene: %d triangles, BVH build success : total %d nodes (%s, %.2f s)\n", nTriangles, nNodes,
                o.sah ? "buildBVHwithSAH" : "buildBVH", now() - t0);

    // ---- environment (M:843-853) ----
    t0 = now();
    int hw = o.hdr_w, hh = o.hdr_h;
    float* hdr = nullptr;
    std::vector<float> synth;
    if (!o.hdr.empty()) {
        ptscheck(pts_hdr_load(o.hdr.c_str(), &hw, &hh, &hdr), "HDRLoader::load");
    } else {
        synth.resize((size_t)hw * hh * 3);
        ptscheck(pts_hdr_synth(hw, hh, 1, synth.data()), "synthetic sky");
    }
    const float* hdr_data = hdr ? hdr : synth.data();
    std::vector<float> cache;
    if (o.variant == PT_VARIANT_DISNEY_IS_MIS) {
        cache.resize((size_t)hw * hh * 3);
        ptscheck(pts_hdr_cache(hdr_data, hw, hh, cache.data()), "calculateHdrCache");
    }
    std::printf("environment: %dx%d%s (%.2f s)\n", hw, hh, cache.empty() ? "" : " + hdrCache", now() - t0);

    float eye[3], cam[16];
    pts_camera(o.up, o.rot, o.r, eye, cam);
    const float* hdr_cache = cache.empty() ? nullptr : cache.data();
    if (o.gpus > 1) {
        run_multi_gpu(o, sc, nTriangles, nNodes, hdr_data, hdr_cache, hw, hh, eye, cam);
        if (hdr) pts_free(hdr);
        pts_scene_destroy(sc);
        return 0;
    }

    // ---- context + upload (M:746-751, M:823-873) ----
    PtConfig cfg = {o.variant, o.width, o.height, o.max_bounce, o.device, o.mode, o.pipeline};
    pt_ctx* pt = nullptr;
    ptcheck(pt_create(&cfg, &pt), "pt_create");
    t0 = now();
    ptcheck(pt_upload_scene(pt, pts_triangles_encoded(sc), nTriangles, pts_nodes_encoded(sc), nNodes, hdr_data, hdr_cache, hw, hh), "pt_upload_scene");
    std::printf("upload: %.2f s\n", now() - t0);

    // ---- display() loop (M:659-709) ----
    ptcheck(pt_set_camera(pt, eye, cam), "pt_set_camera");
    t0 = now();
    unsigned frameCounter = 0;
    while (frameCounter < o.spp) {
        const unsigned n = o.spp - frameCounter < o.per_call ? o.spp - frameCounter : o.per_call;
        ptcheck(pt_render_spp(pt, frameCounter, n), "pt_render_spp");
        frameCounter += n;
    }
    ptcheck(pt_synchronize(pt), "pt_synchronize");
    const double wall = now() - t0;
    PtStats st;
    ptcheck(pt_get_stats(pt, &st), "pt_get_stats");
    std::printf("rendered %u spp at %dx%d: %.2f ms on the device (%.2f ms wall), %.1f Mrays/s, %.1f spp/s, %llu launches\n", o.spp, o.width,
                o.height, st.render_ms, wall * 1e3, (double)st.rays / (st.render_ms * 1e3), o.spp / (st.render_ms * 1e-3),
                (unsigned long long)st.launches);

    // ---- pass3 to a file instead of a window ----
    std::vector<unsigned char> rgba8((size_t)o.width * o.height * 4);
    ptcheck(pt_readback_display(pt, rgba8.data(), rgba8.size()), "pt_readback_display");
    ptscheck(pts_write_png(o.out.c_str(), rgba8.data(), o.width, o.height, 4), "pts_write_png");
    std::printf("wrote %s\n", o.out.c_str());
    if (!o.pfm.empty()) {
        std::vector<float> rgba((size_t)o.width * o.height * 4);
        ptcheck(pt_readback(pt, rgba.data(), rgba.size() * sizeof(float)), "pt_readback");
        ptscheck(pts_write_pfm(o.pfm.c_str(), rgba.data(), o.width, o.height), "pts_write_pfm");
        std::printf("wrote %s\n", o.pfm.c_str());
    }
    pt_destroy(pt);
    if (hdr) pts_free(hdr);
    pts_scene_destroy(sc);
    return 0;
}
