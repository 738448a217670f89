// oracle/_ref/ref_gpu_harness — drives the UNMODIFIED reference gpu-version headers on the host.
//
// TEST INFRASTRUCTURE ONLY.  This file is ours; every hit/scatter/camera/transform routine it calls is
// the reference's own code, included from /root/reference/gpu-version through the host shim in
// oracle/ref_build/shim.  It exists to (1) pin oracle/rt_oracle.c against the
// real reference (golden vectors under tests/golden/ are produced by tests/golden/make_golden.py
// from this binary) and (2) serve as the timed CPU baseline for JSON scenes.
//
// ray_color()/render() live in main.cu next to <<<>>> launch syntax and cannot be included; the two
// loops below restate main.cu:17-70 and main.cu:95-104 statement by statement.
//
// Modes (all binary I/O is raw little-endian float32):
//   aov      <scene.json> <out.bin>              per-pixel first-hit record, 24 floats / pixel
//   render   <scene.json> <seed> <out.bin> [nthreads]   W*H*3 radiance SUMS (row 0 = bottom), prints stats JSON
//   kat_sphere|kat_rect|kat_cyl|kat_scatter|kat_camera|kat_xform <in.bin> <out.bin>
#include "parser.hpp"

#include <cfloat>
#include <chrono>
#include <string>
#include <thread>
#include <vector>
#include <atomic>
#include <unistd.h>
#include <fcntl.h>

// ---- restatement of main.cu:17-70 (ray_color), instrumented with a ray counter ----
static color ref_ray_color(const ray &r, const color &background, hittable **world, int depth,
                           curandState *rng, long long *nrays) {
    ray now = r;
    color accumulated_attenuation(1.0f, 1.0f, 1.0f);
    color accumulated_color(0.0f, 0.0f, 0.0f);
    while (depth > 0) {
        hit_record rec;
        ++*nrays;
        if ((*world)->hit(now, 0.001, FLT_MAX, rec)) {
            ray scattered;
            color attenuation;
            color emitted = rec.mat_ptr->emitted(rec.u, rec.v, rec.p);
            if (rec.mat_ptr->scatter(now, rec, attenuation, scattered, rng)) {
                accumulated_color += emitted * accumulated_attenuation;
                accumulated_attenuation = accumulated_attenuation * attenuation;
                depth--;
                now = scattered;
                continue;
            } else {
                accumulated_color += accumulated_attenuation * emitted;
                break;
            }
        } else {
            accumulated_color += accumulated_attenuation * background;
            break;
        }
    }
    return accumulated_color;
}

static std::vector<float> read_floats(const char *path) {
    FILE *f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
    fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
    std::vector<float> v(n / 4);
    if (fread(v.data(), 4, v.size(), f) != v.size()) { fprintf(stderr, "short read\n"); exit(2); }
    fclose(f);
    return v;
}
static void write_floats(const char *path, const std::vector<float> &v) {
    FILE *f = fopen(path, "wb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
    fwrite(v.data(), 4, v.size(), f);
    fclose(f);
}

// parse_scene dumps the JSON on stdout twice (parser.hpp:508,570): silence it.
static scene *quiet_parse(const std::string &file, json &data) {
    fflush(stdout);
    int saved = dup(1);
    int nul = open("/dev/null", O_WRONLY);
    dup2(nul, 1);
    scene *w;
    std::tie(w, data) = parse_scene(file);
    fflush(stdout);
    std::cout.flush();
    dup2(saved, 1);
    close(nul); close(saved);
    return w;
}

static int index_of(void **arr, int n, void *p) {
    for (int i = 0; i < n; ++i) if (arr[i] == p) return i;
    return -1;
}

static int mode_aov(const char *scene_file, const char *out) {
    json data;
    scene *w = quiet_parse(scene_file, data);
    const int W = w->width, H = w->height;
    camera *cam = *w->cam;
    hittable_list *list = (hittable_list *)(*w->world);
    std::vector<float> o((size_t)W * H * 24, 0.f);
    curandState rng; curand_init(0, 0, 0, &rng);
    static const float fixed_u[8] = {0.5f, 0.5f, 0.5f, 0.25f, 0.25f, 0.25f, 0.1f, 0.1f};
    for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
        float *q = &o[((size_t)y * W + x) * 24];
        auto u = float(x + 0.5f) / (W - 1);        // main.cu:97-98 with the jitter fixed at 0.5
        auto v = float(y + 0.5f) / (H - 1);
        ray r = cam->get_ray(u, v, &rng);
        // hittable_list::hit (object.cuh:23-37), unrolled here to learn the winning index
        hit_record rec, tmp; bool any = false; float closest = FLT_MAX; int prim = -1;
        for (int i = 0; i < list->len; ++i)
            if (list->objects[i]->hit(r, 0.001, closest, tmp)) { any = true; closest = tmp.t; rec = tmp; prim = i; }
        hit_record rec2;
        bool any2 = (*w->world)->hit(r, 0.001, FLT_MAX, rec2);
        if (any != any2 || (any && rec.t != rec2.t)) { fprintf(stderr, "list walk mismatch\n"); return 3; }
        q[0] = any ? 1.f : 0.f;
        if (any) {
            q[1] = rec.t; q[2] = rec.p.x(); q[3] = rec.p.y(); q[4] = rec.p.z();
            q[5] = rec.normal.x(); q[6] = rec.normal.y(); q[7] = rec.normal.z();
            q[8] = rec.front_face ? 1.f : 0.f; q[9] = rec.u; q[10] = rec.v;
            q[11] = (float)prim;
            q[12] = (float)index_of((void **)w->materials, w->num_of_materials, rec.mat_ptr);
            color att(0, 0, 0); ray sc;
            curandState rs; rs.replay = fixed_u; rs.replay_n = 8;
            bool ok = rec.mat_ptr->scatter(r, rec, att, sc, &rs);
            if (!ok && rec.mat_ptr->type == class_type::diffuse_light) att = color(0, 0, 0);
            q[13] = att.x(); q[14] = att.y(); q[15] = att.z();
            color e = rec.mat_ptr->emitted(rec.u, rec.v, rec.p);
            q[16] = e.x(); q[17] = e.y(); q[18] = e.z();
            q[22] = ok ? 1.f : 0.f;
        } else {
            q[19] = w->background.x(); q[20] = w->background.y(); q[21] = w->background.z();
        }
    }
    write_floats(out, o);
    fprintf(stderr, "{\"mode\":\"aov\",\"width\":%d,\"height\":%d,\"objects\":%d}\n", W, H, w->num_of_objects);
    return 0;
}

static int mode_render(const char *scene_file, unsigned long long seed, const char *out, int nthreads) {
    json data;
    scene *w = quiet_parse(scene_file, data);
    const int W = w->width, H = w->height, spp = w->samples_per_pixel;
    camera *cam = *w->cam;
    std::vector<float> img((size_t)W * H * 3);
    std::atomic<int> next_row(0);
    std::vector<long long> rays(nthreads, 0);
    auto t0 = std::chrono::steady_clock::now();
    auto work = [&](int tid) {
        for (;;) {
            int y = next_row.fetch_add(1);
            if (y >= H) break;
            for (int x = 0; x < W; ++x) {
                int id = y * W + x;
                curandState st; curand_init(seed * 0x100000000ULL + id, 0, 0, &st);   // main.cu:120-125: one stream per pixel
                color res(0, 0, 0);
                for (int s = 0; s < spp; ++s) {                                          // main.cu:95-103
                    auto u = float(x + random_float(&st)) / (W - 1);
                    auto v = float(y + random_float(&st)) / (H - 1);
                    ray r = cam->get_ray(u, v, &st);
                    res += ref_ray_color(r, w->background, w->world, w->max_depth, &st, &rays[tid]);
                }
                img[(size_t)id * 3 + 0] = res.x(); img[(size_t)id * 3 + 1] = res.y(); img[(size_t)id * 3 + 2] = res.z();
            }
        }
    };
    std::vector<std::thread> th;
    for (int i = 0; i < nthreads; ++i) th.emplace_back(work, i);
    for (auto &t : th) t.join();
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    long long nr = 0; for (auto r : rays) nr += r;
    write_floats(out, img);
    printf("{\"mode\":\"render\",\"width\":%d,\"height\":%d,\"spp\":%d,\"threads\":%d,\"seconds\":%.6f,"
           "\"samples\":%lld,\"rays\":%lld,\"samples_per_s\":%.1f,\"mrays_per_s\":%.4f}\n",
           W, H, spp, nthreads, sec, (long long)W * H * spp, nr, (double)W * H * spp / sec, nr / sec / 1e6);
    return 0;
}

static void put_rec(float *q, bool hit, const hit_record &rec) {
    q[0] = hit ? 1.f : 0.f;
    if (!hit) return;
    q[1] = rec.t; q[2] = rec.p.x(); q[3] = rec.p.y(); q[4] = rec.p.z();
    q[5] = rec.normal.x(); q[6] = rec.normal.y(); q[7] = rec.normal.z();
    q[8] = rec.front_face ? 1.f : 0.f; q[9] = rec.u; q[10] = rec.v;
}

static int mode_kat(const std::string &kind, const char *in, const char *out) {
    std::vector<float> a = read_floats(in), o;
    lambertian dummy(color(0.5f, 0.5f, 0.5f));
    if (kind == "kat_sphere") {           // in: o3 d3 c3 r tmin tmax (12)  out: 11
        size_t n = a.size() / 12; o.assign(n * 11, 0.f);
        for (size_t i = 0; i < n; ++i) {
            const float *p = &a[i * 12];
            sphere s(point3(p[6], p[7], p[8]), p[9], &dummy);
            hit_record rec; ray r(point3(p[0], p[1], p[2]), vec3(p[3], p[4], p[5]));
            put_rec(&o[i * 11], s.hit(r, p[10], p[11], rec), rec);
        }
    } else if (kind == "kat_rect") {      // in: axis o3 d3 a0 a1 b0 b1 k tmin tmax (14)  out: 11
        size_t n = a.size() / 14; o.assign(n * 11, 0.f);
        for (size_t i = 0; i < n; ++i) {
            const float *p = &a[i * 14];
            hit_record rec; ray r(point3(p[1], p[2], p[3]), vec3(p[4], p[5], p[6]));
            bool h;
            if (p[0] == 0.f) { xy_rect q(p[7], p[8], p[9], p[10], p[11], &dummy); h = q.hit(r, p[12], p[13], rec); }
            else if (p[0] == 1.f) { xz_rect q(p[7], p[8], p[9], p[10], p[11], &dummy); h = q.hit(r, p[12], p[13], rec); }
            else { yz_rect q(p[7], p[8], p[9], p[10], p[11], &dummy); h = q.hit(r, p[12], p[13], rec); }
            put_rec(&o[i * 11], h, rec);
        }
    } else if (kind == "kat_cyl") {       // in: o3 d3 radius zmin zmax axis3 angle_deg tr3 tmin tmax (18)  out: 11
        size_t n = a.size() / 18; o.assign(n * 11, 0.f);
        for (size_t i = 0; i < n; ++i) {
            const float *p = &a[i * 18];
            cylinder c(p[6], p[7], p[8], &dummy);
            c.rotate(vec3(p[9], p[10], p[11]), double(p[12]) / 180.0 * M_PI);     // parser.hpp:423-431
            c.translate(vec3(p[13], p[14], p[15]));                               // parser.hpp:432-438
            hit_record rec; ray r(point3(p[0], p[1], p[2]), vec3(p[3], p[4], p[5]));
            put_rec(&o[i * 11], c.hit(r, p[16], p[17], rec), rec);
        }
    } else if (kind == "kat_scatter") {   // in: type d3 p3 n3 front albedo3 fuzz ior uniforms[16] (32)  out: ok att3 o3 d3 draws (11)
        size_t n = a.size() / 32; o.assign(n * 11, 0.f);
        for (size_t i = 0; i < n; ++i) {
            const float *p = &a[i * 32];
            hit_record rec; rec.p = point3(p[4], p[5], p[6]); rec.normal = vec3(p[7], p[8], p[9]);
            rec.front_face = p[10] != 0.f; rec.t = 1.f; rec.u = 0.f; rec.v = 0.f;
            ray r(point3(0, 0, 0), vec3(p[1], p[2], p[3]));
            color alb(p[11], p[12], p[13]);
            curandState rs; rs.replay = p + 16; rs.replay_n = 16;
            color att(0, 0, 0); ray sc(point3(0, 0, 0), vec3(0, 0, 0)); bool ok;
            int t = (int)p[0];
            if (t == 0) { lambertian m(alb); ok = m.scatter(r, rec, att, sc, &rs); }
            else if (t == 1) { metal m(alb, p[14]); ok = m.scatter(r, rec, att, sc, &rs); }
            else if (t == 2) { dielectric m(p[15]); ok = m.scatter(r, rec, att, sc, &rs); }
            else { diffuse_light m(alb); ok = m.scatter(r, rec, att, sc, &rs); att = m.emitted(0, 0, rec.p); }
            float *q = &o[i * 11];
            q[0] = ok; q[1] = att.x(); q[2] = att.y(); q[3] = att.z();
            q[4] = sc.orig.x(); q[5] = sc.orig.y(); q[6] = sc.orig.z();
            q[7] = sc.dir.x(); q[8] = sc.dir.y(); q[9] = sc.dir.z(); q[10] = (float)rs.draws;
        }
    } else if (kind == "kat_camera") {    // in: from3 at3 vup3 vfov aspect aperture s t (14)  out: o3 d3 (6)
        size_t n = a.size() / 14; o.assign(n * 6, 0.f);
        curandState rs; curand_init(1, 0, 0, &rs);
        for (size_t i = 0; i < n; ++i) {
            const float *p = &a[i * 14];
            point3 from(p[0], p[1], p[2]), at(p[3], p[4], p[5]);
            camera c(from, at, vec3(p[6], p[7], p[8]), p[9], p[10], p[11], (from - at).length());   // parser.hpp:113-126
            ray r = c.get_ray(p[12], p[13], &rs);
            float *q = &o[i * 6];
            q[0] = r.orig.x(); q[1] = r.orig.y(); q[2] = r.orig.z(); q[3] = r.dir.x(); q[4] = r.dir.y(); q[5] = r.dir.z();
        }
    } else if (kind == "kat_xform") {     // in: axis3 angle_deg tr3 p3 (10)  out: point3 vec3 normal3 invpoint3 (12)
        size_t n = a.size() / 10; o.assign(n * 12, 0.f);
        for (size_t i = 0; i < n; ++i) {
            const float *p = &a[i * 10];
            cylinder c(1, -1, 1, &dummy);
            c.rotate(vec3(p[0], p[1], p[2]), double(p[3]) / 180.0 * M_PI);
            c.translate(vec3(p[4], p[5], p[6]));
            vec3 v(p[7], p[8], p[9]);
            vec3 r0 = c.o2w.apply_point(v), r1 = c.o2w.apply_vec(v), r2 = c.o2w.apply_normal(v),
                 r3 = c.o2w.inverse().apply_point(v);
            float *q = &o[i * 12];
            for (int k = 0; k < 3; ++k) { q[k] = r0[k]; q[3 + k] = r1[k]; q[6 + k] = r2[k]; q[9 + k] = r3[k]; }
        }
    } else {
        fprintf(stderr, "unknown kat kind %s\n", kind.c_str());
        return 2;
    }
    write_floats(out, o);
    return 0;
}

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: see header of ref_gpu_harness.cpp\n"); return 2; }
    std::string mode = argv[1];
    if (mode == "aov" && argc >= 4) return mode_aov(argv[2], argv[3]);
    if (mode == "render" && argc >= 5)
        return mode_render(argv[2], strtoull(argv[3], 0, 10), argv[4], argc > 5 ? atoi(argv[5]) : 1);
    if (mode.rfind("kat_", 0) == 0 && argc >= 4) return mode_kat(mode, argv[2], argv[3]);
    fprintf(stderr, "bad arguments\n");
    return 2;
}
