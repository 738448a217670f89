// rt_json.cpp — JSON scene reader: same schema as gpu-version/parser.hpp:504-573 (the reference uses nlohmann::json,
// which is not part of this build), flattened straight into rt_scene_desc tables instead of a pointer graph.
//
// Reference behaviour kept: cross-references are integer indices; unknown keys are ignored (e.g. "id" in blue.json,
// "texture" on a dielectric in sample_scene.json); aspect = width/height and focus distance = |lookfrom - lookat|
// (parser.hpp:113-126); cylinder transform = translate * rotate with the angle in degrees (parser.hpp:423-440);
// "output_file" defaults to "main.png" (parser.hpp:567-568).
// Reference behaviour changed: malformed input returns RT_ERR_PARSE with a message instead of terminating the
// process, and an unknown "type" is an error instead of a silently null slot.
// Extensions (all optional): "background_mode":"sky_gradient", "sampling":"reference_cpu", camera "aspect_ratio" /
// "focus_dist", textures "checker"{even,odd} and "image"{file}, objects "box"{min,max[,rotate,translate]},
// "triangle"{v0,v1,v2} and "mesh"{file(.obj),material[,scale,translate]}.
#include "rt_ctx.h"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>

namespace {

struct JValue;
using JPtr = std::shared_ptr<JValue>;
struct JValue {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    double num = 0; bool b = false; std::string str;
    std::vector<JPtr> arr;
    std::vector<std::pair<std::string, JPtr>> obj;
    const JValue *get(const char *key) const {
        if (kind != Obj) return nullptr;
        const JValue *found = nullptr;
        for (auto &kv : obj) if (kv.first == key) found = kv.second.get();   // last duplicate wins, as nlohmann does
        return found;
    }
};

struct Parser {
    const char *p, *end; std::string err;
    explicit Parser(const char *t) : p(t), end(t + strlen(t)) {}
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p; }
    bool fail(const std::string &m) { if (err.empty()) err = m + " at offset " + std::to_string((long long)(end - p)) + " from end"; return false; }
    bool parse(JPtr &out, int depth = 0) {
        if (depth > 64) return fail("nesting too deep");
        ws();
        if (p >= end) return fail("unexpected end of input");
        out = std::make_shared<JValue>();
        char c = *p;
        if (c == '{') {
            out->kind = JValue::Obj; ++p; ws();
            if (p < end && *p == '}') { ++p; return true; }
            for (;;) {
                ws();
                if (p >= end || *p != '"') return fail("expected string key");
                std::string key; if (!str(key)) return false;
                ws();
                if (p >= end || *p != ':') return fail("expected ':'");
                ++p;
                JPtr v; if (!parse(v, depth + 1)) return false;
                out->obj.emplace_back(key, v);
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == '}') { ++p; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (c == '[') {
            out->kind = JValue::Arr; ++p; ws();
            if (p < end && *p == ']') { ++p; return true; }
            for (;;) {
                JPtr v; if (!parse(v, depth + 1)) return false;
                out->arr.push_back(v);
                ws();
                if (p < end && *p == ',') { ++p; continue; }
                if (p < end && *p == ']') { ++p; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (c == '"') { out->kind = JValue::Str; return str(out->str); }
        if (!strncmp(p, "true", 4)) { out->kind = JValue::Bool; out->b = true; p += 4; return true; }
        if (!strncmp(p, "false", 5)) { out->kind = JValue::Bool; out->b = false; p += 5; return true; }
        if (!strncmp(p, "null", 4)) { out->kind = JValue::Null; p += 4; return true; }
        char *e = nullptr;
        double v = strtod(p, &e);
        if (e == p) return fail("unexpected character");
        out->kind = JValue::Num; out->num = v; p = e;
        return true;
    }
    bool str(std::string &out) {
        ++p;
        while (p < end && *p != '"') {
            if (*p == '\\') {
                ++p; if (p >= end) return fail("bad escape");
                switch (*p) {
                case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
                case 'b': out += '\b'; break; case 'f': out += '\f'; break;
                case 'u': {
                    if (end - p < 5) return fail("bad \\u escape");
                    unsigned cp = (unsigned)strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16); p += 4;
                    if (cp < 0x80) out += (char)cp;
                    else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
                    else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
                    break;
                }
                default: out += *p;
                }
                ++p;
            } else out += *p++;
        }
        if (p >= end) return fail("unterminated string");
        ++p;
        return true;
    }
};

}  // namespace

// translate(t) * rotate(axis, angle) and its inverse: vec3.cuh:382-416 + object.cuh:225-231 + parser.hpp:423-440.  One function for the JSON
// reader and for rt_make_xform (callers that update transforms between frames get bit-identical matrices).
static void mul4(const float a[16], const float b[16], float r[16]) {
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j)
        r[i * 4 + j] = a[i * 4 + 0] * b[0 * 4 + j] + a[i * 4 + 1] * b[1 * 4 + j] + a[i * 4 + 2] * b[2 * 4 + j] + a[i * 4 + 3] * b[3 * 4 + j];
}
void rt_build_xform(const float axis[3], double angle_deg, bool has_rotate, const float t3[3], bool has_translate, rt_xform_desc *out) {
    float m[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, mi[16];
    memcpy(mi, m, sizeof m);
    if (has_rotate) {
        float theta = (float)(angle_deg / 180.0 * M_PI);
        float len = sqrtf(axis[0] * axis[0] + axis[1] * axis[1] + axis[2] * axis[2]);
        float ax = axis[0] / len, ay = axis[1] / len, az = axis[2] / len;
        float s = sinf(theta), c = cosf(theta);
        float r[16] = {ax * ax + (1 - ax * ax) * c, ax * ay * (1 - c) - az * s, ax * az * (1 - c) + ay * s, 0,
                       ax * ay * (1 - c) + az * s, ay * ay + (1 - ay * ay) * c, ay * az * (1 - c) - ax * s, 0,
                       ax * az * (1 - c) - ay * s, ay * az * (1 - c) + ax * s, az * az + (1 - az * az) * c, 0,
                       0, 0, 0, 1};
        float rt[16], nm[16], nmi[16];
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) rt[i * 4 + j] = r[j * 4 + i];
        mul4(r, m, nm); mul4(mi, rt, nmi);
        memcpy(m, nm, sizeof m); memcpy(mi, nmi, sizeof mi);
    }
    if (has_translate) {
        float t[16] = {1, 0, 0, t3[0], 0, 1, 0, t3[1], 0, 0, 1, t3[2], 0, 0, 0, 1};
        float ti[16] = {1, 0, 0, -t3[0], 0, 1, 0, -t3[1], 0, 0, 1, -t3[2], 0, 0, 0, 1};
        float nm[16], nmi[16];
        mul4(t, m, nm); mul4(mi, ti, nmi);
        memcpy(m, nm, sizeof m); memcpy(mi, nmi, sizeof mi);
    }
    memcpy(out->m, m, sizeof out->m); memcpy(out->m_inv, mi, sizeof out->m_inv);
}

namespace {
struct Loader {
    HostScene &hs; std::string base; std::string err;
    bool fail(const std::string &m) { if (err.empty()) err = m; return false; }
    bool num(const JValue *o, const char *key, double &out, const std::string &where) {
        const JValue *v = o ? o->get(key) : nullptr;
        if (!v || v->kind != JValue::Num) return fail(where + ": missing number \"" + key + "\"");
        out = v->num; return true;
    }
    bool vec(const JValue *o, const char *key, int n, float *out, const std::string &where) {
        const JValue *v = o ? o->get(key) : nullptr;
        if (!v || v->kind != JValue::Arr || (int)v->arr.size() < n) return fail(where + ": missing array \"" + key + "\"");
        for (int i = 0; i < n; ++i) { if (v->arr[i]->kind != JValue::Num) return fail(where + ": \"" + key + "\" must hold numbers"); out[i] = (float)v->arr[i]->num; }
        return true;
    }
    bool xform(const JValue *o, const std::string &where, int &index, bool always) {
        const JValue *rot = o->get("rotate"), *tr = o->get("translate");
        if (!rot && !tr && !always) { index = -1; return true; }
        float axis[3] = {0, 0, 1}, t3[3] = {0, 0, 0}; double angle = 0.0;
        if (rot && (!vec(rot, "axis", 3, axis, where + ".rotate") || !num(rot, "angle", angle, where + ".rotate"))) return false;
        if (tr && !vec(o, "translate", 3, t3, where)) return false;
        rt_xform_desc x;
        rt_build_xform(axis, angle, rot != nullptr, t3, tr != nullptr, &x);
        index = (int)hs.xforms.size();
        hs.xforms.push_back(x);
        return true;
    }
    // p <- m * p for a 3x4 row-major matrix, plain left-to-right float arithmetic (the Python loader of the tests states the same)
    static void bake(const float *m, float *p) {
        const float x = p[0], y = p[1], z = p[2];
        for (int r = 0; r < 3; ++r) p[r] = ((m[4 * r] * x + m[4 * r + 1] * y) + m[4 * r + 2] * z) + m[4 * r + 3];
    }
    // (u, v) clamped to [0, 1] as two 16-bit unorm values in the bits of one float (rt.h: TRIANGLE p[9..11])
    static float pack_uv(float u, float v) {
        u = u < 0.0f ? 0.0f : (u > 1.0f ? 1.0f : u); v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
        const uint32_t q = (uint32_t)(u * 65535.0f + 0.5f) | ((uint32_t)(v * 65535.0f + 0.5f) << 16);
        float f; memcpy(&f, &q, 4);
        return f;
    }
    std::string resolve(const std::string &f) const { return (!f.empty() && f[0] == '/') || base.empty() ? f : base + "/" + f; }
    // Wavefront OBJ: v, vt and f (v, v/vt, v/vt/vn, v//vn; negative indices; polygons fan-triangulated).  With `rot` (a 3x4 matrix whose
    // translation column is the mesh's translate) a vertex is m * (scale * v), else scale * v + tr.  `vt` -> per-vertex uv of the triangles.
    bool load_obj(const std::string &file, int material, float scale, const float tr[3], const float *rot, const std::string &where) {
        std::ifstream in(resolve(file));
        if (!in) return fail(where + ": cannot open mesh file " + file);
        std::vector<float> v, vt; std::string line;
        while (std::getline(in, line)) {
            std::istringstream ls(line); std::string tag; ls >> tag;
            if (tag == "v") {
                float p[3]; ls >> p[0] >> p[1] >> p[2];
                if (rot) { for (int a = 0; a < 3; ++a) p[a] = p[a] * scale; bake(rot, p); }
                else for (int a = 0; a < 3; ++a) p[a] = p[a] * scale + tr[a];
                v.insert(v.end(), p, p + 3);
            } else if (tag == "vt") { float a = 0, b = 0; ls >> a >> b; vt.push_back(a); vt.push_back(b); }
            else if (tag == "f") {
                std::vector<int> idx, tidx; std::string tok;
                while (ls >> tok) {
                    int i = atoi(tok.c_str());
                    idx.push_back(i < 0 ? (int)(v.size() / 3) + i : i - 1);
                    int ti = -1;
                    const size_t sl = tok.find('/');
                    if (sl != std::string::npos && sl + 1 < tok.size() && tok[sl + 1] != '/') {
                        const int j = atoi(tok.c_str() + sl + 1);
                        ti = j < 0 ? (int)(vt.size() / 2) + j : j - 1;
                    }
                    tidx.push_back(ti);
                }
                for (size_t k = 1; k + 1 < idx.size(); ++k) {     // fan triangulation
                    rt_object_desc o{}; o.type = RT_OBJ_TRIANGLE; o.material = material; o.xform = -1;
                    const int tri[3] = {idx[0], idx[k], idx[k + 1]}, tt[3] = {tidx[0], tidx[k], tidx[k + 1]};
                    const bool has_uv = tt[0] >= 0 && tt[1] >= 0 && tt[2] >= 0;
                    for (int c = 0; c < 3; ++c) {
                        if (tri[c] < 0 || (size_t)tri[c] * 3 + 2 >= v.size()) return fail(where + ": face index out of range in " + file);
                        for (int a = 0; a < 3; ++a) o.p[3 * c + a] = v[3 * tri[c] + a];
                        if (has_uv) {
                            if ((size_t)tt[c] * 2 + 1 >= vt.size()) return fail(where + ": texture index out of range in " + file);
                            o.p[9 + c] = pack_uv(vt[2 * tt[c]], vt[2 * tt[c] + 1]);
                        }
                    }
                    hs.objects.push_back(o);
                }
            }
        }
        return true;
    }
    bool run(const JValue &root) {
        rt_scene_desc &d = hs.d;
        double x;
        float bg[3];
        if (!vec(&root, "background", 3, bg, "scene")) return false;
        memcpy(d.background, bg, sizeof bg);
        if (!num(&root, "max_depth", x, "scene")) return false; d.max_depth = (int)x;
        if (!num(&root, "samples_per_pixel", x, "scene")) return false; d.samples_per_pixel = (int)x;
        if (!num(&root, "width", x, "scene")) return false; d.width = (int)x;
        if (!num(&root, "height", x, "scene")) return false; d.height = (int)x;
        const JValue *v = root.get("output_file");
        if (v && v->kind == JValue::Str) hs.output_file = v->str;
        v = root.get("background_mode");
        d.background_mode = (v && v->kind == JValue::Str && v->str == "sky_gradient") ? RT_BG_SKY_GRADIENT : RT_BG_CONSTANT;
        v = root.get("sampling");
        d.sampling = (v && v->kind == JValue::Str && v->str == "reference_cpu") ? RT_SAMPLING_REFERENCE_CPU : RT_SAMPLING_REFERENCE_GPU;
        v = root.get("sampler");          // "rejection": the reference's loops; default: the same distributions drawn directly
        if (!(v && v->kind == JValue::Str && v->str == "rejection")) d.sampling |= RT_SAMPLING_DIRECT;
        // camera, parser.hpp:113-126
        const JValue *cam = root.get("camera");
        if (!cam) return fail("scene: missing \"camera\"");
        rt_camera_desc &c = d.camera;
        if (!vec(cam, "lookfrom", 3, c.lookfrom, "camera") || !vec(cam, "lookat", 3, c.lookat, "camera") || !vec(cam, "vup", 3, c.vup, "camera")) return false;
        if (!num(cam, "vfov", x, "camera")) return false; c.vfov_deg = (float)x;
        if (!num(cam, "aperture", x, "camera")) return false; c.aperture = (float)x;
        c.aspect = (float)d.width / (float)d.height;
        const JValue *ar = cam->get("aspect_ratio");
        if (ar && ar->kind == JValue::Num) c.aspect = (float)ar->num;
        float dx = c.lookfrom[0] - c.lookat[0], dy = c.lookfrom[1] - c.lookat[1], dz = c.lookfrom[2] - c.lookat[2];
        c.focus_dist = sqrtf(dx * dx + dy * dy + dz * dz);
        const JValue *fd = cam->get("focus_dist");
        if (fd && fd->kind == JValue::Num) c.focus_dist = (float)fd->num;
        auto table = [&](const char *name) -> const JValue * {
            const JValue *t = root.get(name);
            const JValue *dd = t ? t->get("data") : nullptr;
            return (dd && dd->kind == JValue::Arr) ? dd : nullptr;
        };
        const JValue *texs = table("texture"), *mats = table("material"), *objs = table("object");
        if (!texs || !mats || !objs) return fail("scene: \"texture\", \"material\" and \"object\" must each hold a \"data\" array");
        for (size_t i = 0; i < texs->arr.size(); ++i) {      // parser.hpp:143-184
            const JValue *t = texs->arr[i].get(); const std::string where = "texture[" + std::to_string(i) + "]";
            const JValue *ty = t->get("type");
            if (!ty || ty->kind != JValue::Str) return fail(where + ": missing \"type\"");
            rt_texture_desc q{};
            if (ty->str == "solid_color") { q.type = RT_TEX_SOLID; if (!vec(t, "color", 3, q.color, where)) return false; }
            else if (ty->str == "checker") {
                q.type = RT_TEX_CHECKER;
                if (!num(t, "even", x, where)) return false; q.even = (int)x;
                if (!num(t, "odd", x, where)) return false; q.odd = (int)x;
            } else if (ty->str == "image") {
                q.type = RT_TEX_IMAGE;
                const JValue *f = t->get("file");
                if (!f || f->kind != JValue::Str) return fail(where + ": missing \"file\"");
                std::vector<uint8_t> rgb; int w = 0, h = 0; std::string e;
                if (rt_png_decode_file(resolve(f->str).c_str(), rgb, w, h, e)) return fail(where + ": " + e);
                q.image = (int)hs.images.size();
                rt_image_desc im{}; im.width = w; im.height = h;
                hs.images.push_back(im); hs.image_data.push_back(std::move(rgb));
            } else return fail(where + ": unknown texture type \"" + ty->str + "\"");
            hs.textures.push_back(q);
        }
        for (size_t i = 0; i < mats->arr.size(); ++i) {      // parser.hpp:186-281
            const JValue *m = mats->arr[i].get(); const std::string where = "material[" + std::to_string(i) + "]";
            const JValue *ty = m->get("type");
            if (!ty || ty->kind != JValue::Str) return fail(where + ": missing \"type\"");
            rt_material_desc q{};
            if (ty->str == "lambertian") { q.type = RT_MAT_LAMBERTIAN; if (!num(m, "texture", x, where)) return false; q.texture = (int)x; }
            else if (ty->str == "metal") {
                q.type = RT_MAT_METAL;
                if (!vec(m, "albedo", 3, q.albedo, where) || !num(m, "fuzz", x, where)) return false;
                q.fuzz = (float)x;
            } else if (ty->str == "dielectric") { q.type = RT_MAT_DIELECTRIC; if (!num(m, "index_of_refraction", x, where)) return false; q.ior = (float)x; }
            else if (ty->str == "diffuse_light") { q.type = RT_MAT_DIFFUSE_LIGHT; if (!num(m, "texture", x, where)) return false; q.texture = (int)x; }
            else return fail(where + ": unknown material type \"" + ty->str + "\"");
            hs.materials.push_back(q);
        }
        for (size_t i = 0; i < objs->arr.size(); ++i) {      // parser.hpp:283-478
            const JValue *o = objs->arr[i].get(); const std::string where = "object[" + std::to_string(i) + "]";
            const JValue *ty = o->get("type");
            if (!ty || ty->kind != JValue::Str) return fail(where + ": missing \"type\"");
            rt_object_desc q{}; q.xform = -1;
            double mat = 0;
            if (!num(o, "material", mat, where)) return false;
            q.material = (int)mat;
            const std::string &t = ty->str;
            auto five = [&](const char *a0, const char *a1, const char *b0, const char *b1) {
                const char *keys[5] = {a0, a1, b0, b1, "k"};
                for (int k = 0; k < 5; ++k) { double vv; if (!num(o, keys[k], vv, where)) return false; q.p[k] = (float)vv; }
                return true;
            };
            // "rotate" / "translate" are honoured on EVERY object type (the reference's transform machinery is general,
            // vec3.cuh:303-427, although its parser only feeds it for cylinders, parser.hpp:423-440): sphere / rect / box / cylinder
            // as an instance transform, triangle / mesh baked into the vertices
            if (t == "sphere") {
                q.type = RT_OBJ_SPHERE;
                if (!vec(o, "center", 3, q.p, where) || !num(o, "radius", x, where)) return false;
                q.p[3] = (float)x;
                if (!xform(o, where, q.xform, false)) return false;
            }
            else if (t == "xy_rect") { q.type = RT_OBJ_XY_RECT; if (!five("x0", "x1", "y0", "y1") || !xform(o, where, q.xform, false)) return false; }
            else if (t == "xz_rect") { q.type = RT_OBJ_XZ_RECT; if (!five("x0", "x1", "z0", "z1") || !xform(o, where, q.xform, false)) return false; }
            else if (t == "yz_rect") { q.type = RT_OBJ_YZ_RECT; if (!five("y0", "y1", "z0", "z1") || !xform(o, where, q.xform, false)) return false; }
            else if (t == "cylinder") {
                q.type = RT_OBJ_CYLINDER;
                const char *keys[3] = {"radius", "zmin", "zmax"};
                for (int k = 0; k < 3; ++k) { if (!num(o, keys[k], x, where)) return false; q.p[k] = (float)x; }
                if (!xform(o, where, q.xform, true)) return false;
            } else if (t == "box") {
                q.type = RT_OBJ_BOX;
                if (!vec(o, "min", 3, q.p, where) || !vec(o, "max", 3, q.p + 3, where) || !xform(o, where, q.xform, false)) return false;
            } else if (t == "triangle") {
                q.type = RT_OBJ_TRIANGLE;
                if (!vec(o, "v0", 3, q.p, where) || !vec(o, "v1", 3, q.p + 3, where) || !vec(o, "v2", 3, q.p + 6, where)) return false;
                int xi = -1;
                if (!xform(o, where, xi, false)) return false;
                if (xi >= 0) { for (int c = 0; c < 3; ++c) bake(hs.xforms[xi].m, q.p + 3 * c); hs.xforms.pop_back(); }
                const JValue *uvs = o->get("uv");       // optional [[u0,v0],[u1,v1],[u2,v2]]
                if (uvs && uvs->kind == JValue::Arr && uvs->arr.size() >= 3)
                    for (int c = 0; c < 3; ++c) {
                        const JValue *e = uvs->arr[c].get();
                        if (e->kind != JValue::Arr || e->arr.size() < 2 || e->arr[0]->kind != JValue::Num || e->arr[1]->kind != JValue::Num) return fail(where + ": \"uv\" must hold three [u, v] pairs");
                        q.p[9 + c] = pack_uv((float)e->arr[0]->num, (float)e->arr[1]->num);
                    }
            } else if (t == "mesh") {
                const JValue *f = o->get("file");
                if (!f || f->kind != JValue::Str) return fail(where + ": missing \"file\"");
                float scale = 1.0f, tr[3] = {0, 0, 0};
                const JValue *sv = o->get("scale"); if (sv && sv->kind == JValue::Num) scale = (float)sv->num;
                if (o->get("translate") && !vec(o, "translate", 3, tr, where)) return false;
                // "rotate" on a mesh: vertices are scaled, rotated about the origin, then translated (translate(t) * rotate, as for every object)
                int xi = -1;
                const float *rot = nullptr;
                if (o->get("rotate")) {
                    if (!xform(o, where, xi, false)) return false;
                    rot = hs.xforms[xi].m;
                }
                const bool ok = load_obj(f->str, q.material, scale, tr, rot, where);
                if (xi >= 0) hs.xforms.pop_back();
                if (!ok) return false;
                continue;
            } else return fail(where + ": unknown object type \"" + t + "\"");
            hs.objects.push_back(q);
        }
        return true;
    }
};

}  // namespace

int rt_json_to_scene(const char *text, const char *base_dir, HostScene &out, std::string &err) {
    Parser ps(text);
    JPtr root;
    if (!ps.parse(root)) { err = "JSON: " + ps.err; return RT_ERR_PARSE; }
    ps.ws();
    if (ps.p != ps.end) { err = "JSON: trailing characters after the document"; return RT_ERR_PARSE; }
    if (root->kind != JValue::Obj) { err = "JSON: top level must be an object"; return RT_ERR_PARSE; }
    Loader ld{out, base_dir ? base_dir : "", ""};
    if (!ld.run(*root)) { err = ld.err; return RT_ERR_PARSE; }
    return RT_OK;
}
