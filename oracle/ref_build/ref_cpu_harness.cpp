// oracle/_ref/ref_cpu_harness — drives the UNMODIFIED reference cmake-cpu-version sources.
//
// TEST INFRASTRUCTURE ONLY.  The reference's main.cpp (ray_color, render, random_scene, camera,
// sphere::hit, materials — double precision) is #included as-is with `main` renamed; the recipe needs
// -DMT_RANDOM_GENERATOR and <cstring>.  `#define private public` only opens the
// reference's classes for reading (solid_color::color_value, camera members); no behaviour changes.
//
// Modes (raw little-endian float32 outputs):
//   dump_scene <out.json>                  random_scene() under mt19937(5489) written in OUR JSON schema
//   aov <W> <H> <out.bin>                  first-hit record per pixel (24 floats), pixel-centre rays, aperture 0
//   render <W> <H> <spp> <depth> <skip> <out.bin>   reference render() per pixel in the reference's scan order;
//                                          <skip> random_double() draws are discarded after scene generation
//                                          (gives an independent second estimate for tolerance calibration)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <limits>
#include <memory>
#include <random>
#include <vector>
#include <string>
#include <chrono>

#define private public
#define main ref_main
#include "main.cpp"
#undef main
#undef private

static void write_floats(const char *path, const std::vector<float> &v) {
    FILE *f = fopen(path, "wb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); exit(2); }
    fwrite(v.data(), 4, v.size(), f);
    fclose(f);
}

static color tex_solid(texture *t) {
    auto *s = dynamic_cast<solid_color *>(t);
    if (!s) { fprintf(stderr, "unexpected texture type\n"); exit(3); }
    return s->color_value;
}

static int mode_dump(const char *out) {
    hittable_list world = random_scene();
    FILE *f = fopen(out, "w");
    if (!f) return 2;
    // Our schema (superset of gpu-version/parser.hpp's): one texture+material per sphere, in object order.
    std::string tex, mat, obj;
    char buf[512];
    int ntex = 0, nmat = 0;
    for (size_t i = 0; i < world.objects.size(); ++i) {
        const sphere &s = world.objects[i];
        if (auto *l = dynamic_cast<lambertian *>(s.mat_ptr)) {
            if (auto *c = dynamic_cast<checker_texture *>(l->albedo)) {
                color e = tex_solid(c->even.get()), o = tex_solid(c->odd.get());
                snprintf(buf, sizeof buf, "%s{\"type\":\"solid_color\",\"color\":[%.17g,%.17g,%.17g]},"
                         "{\"type\":\"solid_color\",\"color\":[%.17g,%.17g,%.17g]},"
                         "{\"type\":\"checker\",\"even\":%d,\"odd\":%d}", ntex ? "," : "", e.x(), e.y(), e.z(),
                         o.x(), o.y(), o.z(), ntex, ntex + 1);
                tex += buf; ntex += 3;
            } else {
                color a = tex_solid(l->albedo);
                snprintf(buf, sizeof buf, "%s{\"type\":\"solid_color\",\"color\":[%.17g,%.17g,%.17g]}", ntex ? "," : "",
                         a.x(), a.y(), a.z());
                tex += buf; ntex += 1;
            }
            snprintf(buf, sizeof buf, "%s{\"type\":\"lambertian\",\"texture\":%d}", nmat ? "," : "", ntex - 1);
        } else if (auto *m = dynamic_cast<metal *>(s.mat_ptr)) {
            snprintf(buf, sizeof buf, "%s{\"type\":\"metal\",\"albedo\":[%.17g,%.17g,%.17g],\"fuzz\":%.17g}", nmat ? "," : "",
                     m->albedo.x(), m->albedo.y(), m->albedo.z(), m->fuzz);
        } else if (auto *d = dynamic_cast<dielectric *>(s.mat_ptr)) {
            snprintf(buf, sizeof buf, "%s{\"type\":\"dielectric\",\"index_of_refraction\":%.17g}", nmat ? "," : "", d->ir);
        } else { fprintf(stderr, "unexpected material\n"); return 3; }
        mat += buf;
        snprintf(buf, sizeof buf, "%s{\"type\":\"sphere\",\"center\":[%.17g,%.17g,%.17g],\"radius\":%.17g,\"material\":%d}",
                 i ? "," : "", s.center.x(), s.center.y(), s.center.z(), s.radius, nmat);
        obj += buf; nmat += 1;
    }
    // camera per cmake-cpu-version/main.cpp:64,89-94 (aspect pinned to 16/9 regardless of -w/-h)
    fprintf(f, "{\"output_file\":\"rtiow.png\",\"background\":[0.5,0.7,1.0],\"background_mode\":\"sky_gradient\","
               "\"sampling\":\"reference_cpu\",\"max_depth\":50,\"samples_per_pixel\":10,\"width\":1200,\"height\":800,"
               "\"camera\":{\"lookfrom\":[13,2,3],\"lookat\":[0,0,0],\"vup\":[0,1,0],\"vfov\":20,\"aperture\":0.1,"
               "\"aspect_ratio\":%.17g},\n", 16.0 / 9.0);
    fprintf(f, "\"object\":{\"data\":[%s]},\n\"material\":{\"data\":[%s]},\n\"texture\":{\"data\":[%s]}}\n", obj.c_str(),
            mat.c_str(), tex.c_str());
    fclose(f);
    fprintf(stderr, "{\"mode\":\"dump_scene\",\"spheres\":%zu}\n", world.objects.size());
    return 0;
}

static camera make_camera(double aperture) {       // main.cpp:89-94
    point3 lookfrom(13, 2, 3), lookat(0, 0, 0);
    vec3 vup(0, 1, 0);
    auto dist_to_focus = (lookfrom - lookat).length();
    return camera(lookfrom, lookat, vup, 20, 16.0 / 9.0, aperture, dist_to_focus);
}

static int mode_aov(int W, int H, const char *out) {
    hittable_list world = random_scene();
    camera cam = make_camera(0.0);
    std::vector<float> o((size_t)W * H * 24, 0.f);
    for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
        float *q = &o[((size_t)y * W + x) * 24];
        auto u = double(x + 0.5) / (W - 1);
        auto v = double(y + 0.5) / (H - 1);
        ray r = cam.get_ray(u, v);
        hit_record rec, tmp; bool any = false; double closest = inf; int prim = -1;
        for (size_t i = 0; i < world.objects.size(); ++i)        // hittable_list.h:23-37 unrolled to learn the index
            if (world.objects[i].hit(r, 0.001, closest, tmp)) { any = true; closest = tmp.t; rec = tmp; prim = (int)i; }
        q[0] = any;
        if (any) {
            q[1] = rec.t; q[2] = rec.p.x(); q[3] = rec.p.y(); q[4] = rec.p.z();
            q[5] = rec.normal.x(); q[6] = rec.normal.y(); q[7] = rec.normal.z();
            q[8] = rec.front_face; q[9] = rec.u; q[10] = rec.v; q[11] = (float)prim; q[12] = (float)prim;
            color att; ray sc;
            bool ok = rec.mat_ptr->scatter(r, rec, att, sc);
            q[13] = att.x(); q[14] = att.y(); q[15] = att.z(); q[22] = ok;
        } else {
            vec3 unit_dir = unit_vector(r.direction());          // main.cpp:36-38
            auto t = 0.5 * (unit_dir.y() + 1.0);
            color c = (1.0 - t) * color(1.0, 1, 1) + t * color(0.5, 0.7, 1.0);
            q[19] = c.x(); q[20] = c.y(); q[21] = c.z();
        }
    }
    write_floats(out, o);
    return 0;
}

static int mode_render(int W, int H, int spp, int depth, long long skip, const char *out) {
    hittable_list world = random_scene();
    for (long long i = 0; i < skip; ++i) random_double();
    camera cam = make_camera(0.1);
    std::vector<float> img((size_t)W * H * 3);
    auto t0 = std::chrono::steady_clock::now();
    for (int j = H - 1; j >= 0; --j)                             // main.cpp:99-106 scan order
        for (int i = 0; i < W; ++i) {
            color c = render(i, j, spp, cam, world, depth, W, H);
            size_t id = (size_t)j * W + i;
            img[id * 3] = c.x(); img[id * 3 + 1] = c.y(); img[id * 3 + 2] = c.z();
        }
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    write_floats(out, img);
    printf("{\"mode\":\"render\",\"width\":%d,\"height\":%d,\"spp\":%d,\"threads\":1,\"seconds\":%.6f,"
           "\"samples\":%lld,\"samples_per_s\":%.1f}\n", W, H, spp, sec, (long long)W * H * spp, (double)W * H * spp / sec);
    return 0;
}

int main(int argc, char **argv) {
    std::string mode = argc > 1 ? argv[1] : "";
    if (mode == "dump_scene" && argc >= 3) return mode_dump(argv[2]);
    if (mode == "aov" && argc >= 5) return mode_aov(atoi(argv[2]), atoi(argv[3]), argv[4]);
    if (mode == "render" && argc >= 8)
        return mode_render(atoi(argv[2]), atoi(argv[3]), atoi(argv[4]), atoi(argv[5]), atoll(argv[6]), argv[7]);
    fprintf(stderr, "usage: see header of ref_cpu_harness.cpp\n");
    return 2;
}
