// rt_gltf.cpp — glTF 2.0 / GLB ingestion into the `shared` crate's arrays.
//
// Restates what the reference's loader does with a file (src/gltf_loader.rs:53-125
// extract_scene, :188-226 process_node, :228-287 cameras / KHR_lights_punctual, :290-394
// process_primitive with bit-exact position de-duplication per primitive, :397-489
// convert_material incl. the KHR_materials_{transmission,ior,specular,volume} and
// pbrSpecularGlossiness extensions, :497-590 accessor reads) and src/scene.rs:43-69 (first
// camera or Camera::new, then BvhBuilder::build).  The reference delegates parsing to the
// `gltf` 1.4 and `image` crates, which are not under /root/reference; this file carries its own
// JSON / base64 / GLB reader.  Matrix arithmetic follows glam's scalar-math association
// (Mat4 * Mat4 column by column, transform_point3 = ((X*x + Y*y) + Z*z) + W) in fp32 with
// contraction off; node TRS composition follows the gltf crate's T * R * S.  Parity of the
// loader against the reference is UNPINNED (no fixtures, no tests, crates absent).
//
// Image payloads (PNG / JPEG) are not decoded: TextureInfo records carry the dimensions
// parsed from the image header and a zero-filled RGBA8 block of the right size, which is
// all the path needs — the shader binds textures and never reads them (shader/src/lib.rs:34-35).
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <exception>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/rt_shared.h"
#include "../../include/rtb200_host.h"

namespace {

// ------------------------------------------------------------------ JSON
struct JVal {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<JVal> arr;
    std::vector<std::pair<std::string, JVal>> obj;
    const JVal* get(const char* key) const {
        if (kind != Obj) return nullptr;
        for (auto& kv : obj) if (kv.first == key) return &kv.second;
        return nullptr;
    }
    bool is_num() const { return kind == Num; }
};

struct JParser {
    const char* p; const char* end; std::string err;
    void ws() { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++; }
    bool fail(const char* m) { if (err.empty()) err = m; return false; }
    bool parse_string(std::string& out) {
        if (p >= end || *p != '"') return fail("expected string");
        p++;
        while (p < end && *p != '"') {
            if (*p == '\\') {
                if (++p >= end) return fail("bad escape");
                char c = *p++;
                switch (c) {
                    case '"': out += '"'; break; case '\\': out += '\\'; break; case '/': out += '/'; break;
                    case 'b': out += '\b'; break; case 'f': out += '\f'; break; case 'n': out += '\n'; break;
                    case 'r': out += '\r'; break; case 't': out += '\t'; break;
                    case 'u': {
                        if (end - p < 4) return fail("bad \\u escape");
                        unsigned cp = 0;
                        for (int i = 0; i < 4; i++) {
                            char h = *p++; cp <<= 4;
                            if (h >= '0' && h <= '9') cp |= h - '0'; else if (h >= 'a' && h <= 'f') cp |= h - 'a' + 10;
                            else if (h >= 'A' && h <= 'F') cp |= h - 'A' + 10; else return fail("bad \\u escape");
                        }
                        if (cp < 0x80) out += (char)cp;
                        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
                        else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
                        break;
                    }
                    default: return fail("bad escape");
                }
            } else out += *p++;
        }
        if (p >= end) return fail("unterminated string");
        p++;
        return true;
    }
    bool parse(JVal& v, int depth = 0) {
        if (depth > 200) return fail("JSON nested too deeply");
        ws();
        if (p >= end) return fail("unexpected end of JSON");
        char c = *p;
        if (c == '{') {
            v.kind = JVal::Obj; p++; ws();
            if (p < end && *p == '}') { p++; return true; }
            for (;;) {
                ws();
                std::string k;
                if (!parse_string(k)) return false;
                ws();
                if (p >= end || *p != ':') return fail("expected ':'");
                p++;
                v.obj.emplace_back(std::move(k), JVal());
                if (!parse(v.obj.back().second, depth + 1)) return false;
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == '}') { p++; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (c == '[') {
            v.kind = JVal::Arr; p++; ws();
            if (p < end && *p == ']') { p++; return true; }
            for (;;) {
                v.arr.emplace_back();
                if (!parse(v.arr.back(), depth + 1)) return false;
                ws();
                if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == ']') { p++; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (c == '"') { v.kind = JVal::Str; return parse_string(v.str); }
        if (end - p >= 4 && !strncmp(p, "true", 4)) { v.kind = JVal::Bool; v.b = true; p += 4; return true; }
        if (end - p >= 5 && !strncmp(p, "false", 5)) { v.kind = JVal::Bool; v.b = false; p += 5; return true; }
        if (end - p >= 4 && !strncmp(p, "null", 4)) { v.kind = JVal::Null; p += 4; return true; }
        {
            std::string tmp;
            const char* q = p;
            while (q < end && (strchr("+-.eE", *q) || (*q >= '0' && *q <= '9'))) q++;
            if (q == p) return fail("unexpected character in JSON");
            tmp.assign(p, q);
            char* e2 = nullptr;
            v.num = strtod(tmp.c_str(), &e2);
            if (e2 == tmp.c_str()) return fail("bad number");
            v.kind = JVal::Num; p = q;
            return true;
        }
    }
};

// ------------------------------------------------------------------ helpers
bool base64_decode(const char* s, size_t n, std::vector<uint8_t>& out) {
    static int8_t T[256]; static bool init = false;
    if (!init) {
        memset(T, -1, sizeof T);
        const char* A = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
        for (int i = 0; i < 64; i++) T[(unsigned char)A[i]] = (int8_t)i;
        init = true;
    }
    uint32_t acc = 0; int bits = 0;
    for (size_t i = 0; i < n; i++) {
        unsigned char c = (unsigned char)s[i];
        if (c == '=' || c == '\n' || c == '\r') continue;
        if (T[c] < 0) return false;
        acc = (acc << 6) | (uint32_t)T[c]; bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((uint8_t)(acc >> bits)); }
    }
    return true;
}

bool read_file(const std::string& path, std::vector<uint8_t>& out) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    if (n < 0) { fclose(f); return false; }
    out.resize((size_t)n);
    size_t got = n ? fread(out.data(), 1, (size_t)n, f) : 0;
    fclose(f);
    return got == (size_t)n;
}

// glam Mat4 (column-major), scalar-math association
struct M4 { float c[4][4]; };
M4 m4_identity() { M4 m; memset(&m, 0, sizeof m); for (int i = 0; i < 4; i++) m.c[i][i] = 1.0f; return m; }
// Mat4::mul_vec4: ((x_axis*x + y_axis*y) + z_axis*z) + w_axis*w
void m4_mul_vec4(const M4& a, const float v[4], float out[4]) {
    for (int r = 0; r < 4; r++) {
        float res = a.c[0][r] * v[0];
        res = res + a.c[1][r] * v[1];
        res = res + a.c[2][r] * v[2];
        res = res + a.c[3][r] * v[3];
        out[r] = res;
    }
}
M4 m4_mul(const M4& a, const M4& b) { // Mat4 * Mat4: column j = a * b.col(j)
    M4 o;
    for (int j = 0; j < 4; j++) m4_mul_vec4(a, b.c[j], o.c[j]);
    return o;
}
void m4_point(const M4& m, const float p[3], float out[3]) { // transform_point3
    for (int r = 0; r < 3; r++) {
        float res = m.c[0][r] * p[0];
        res = m.c[1][r] * p[1] + res;
        res = m.c[2][r] * p[2] + res;
        res = m.c[3][r] + res;
        out[r] = res;
    }
}
void m4_vector(const M4& m, const float p[3], float out[3]) { // transform_vector3
    for (int r = 0; r < 3; r++) {
        float res = m.c[0][r] * p[0];
        res = m.c[1][r] * p[1] + res;
        res = m.c[2][r] * p[2] + res;
        out[r] = res;
    }
}
void normalize(float v[3]) { // glam Vec3::normalize = v * (1 / length)
    float len = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    float inv = 1.0f / len;
    v[0] *= inv; v[1] *= inv; v[2] *= inv;
}

struct PosKey {
    uint32_t b[3];
    bool operator==(const PosKey& o) const { return b[0] == o.b[0] && b[1] == o.b[1] && b[2] == o.b[2]; }
};
struct PosHash {
    size_t operator()(const PosKey& k) const {
        uint64_t h = 1469598103934665603ull;
        for (int i = 0; i < 3; i++) { h ^= k.b[i]; h *= 1099511628211ull; }
        return (size_t)h;
    }
};

} // namespace

struct RtGltfScene {
    std::vector<RtTriangle> triangles;
    std::vector<RtVertex> vertices;
    std::vector<RtMaterial> materials;
    std::vector<RtSphere> spheres;
    std::vector<RtLight> lights;
    std::vector<RtCamera> cameras;
    std::vector<RtTextureInfo> textures;
    std::vector<uint8_t> texture_data;
    uint32_t scene_count = 0;
    std::vector<std::string> warnings;
};

namespace {

struct Loader {
    JVal doc;
    std::vector<std::vector<uint8_t>> buffers;
    std::string base_dir, err;
    RtGltfScene* out = nullptr;
    std::map<size_t, uint32_t> material_map;

    bool fail(const std::string& m) { if (err.empty()) err = m; return false; }
    const JVal* arr(const char* key) const { const JVal* v = doc.get(key); return (v && v->kind == JVal::Arr) ? v : nullptr; }
    static bool get_index(const JVal* o, const char* key, size_t* idx) {
        const JVal* v = o ? o->get(key) : nullptr;
        // file-controlled number: NaN, negative and anything past 2^53 (not an exact integer in a JSON
        // double, and large enough to overflow offset arithmetic) is "absent / invalid"
        if (!v || !v->is_num() || !(v->num >= 0.0) || v->num > 9007199254740992.0) return false;
        *idx = (size_t)v->num;
        return true;
    }
    static float num_or(const JVal* o, const char* key, float dflt) {
        const JVal* v = o ? o->get(key) : nullptr;
        return (v && v->is_num()) ? (float)v->num : dflt;
    }
    static void vec_or(const JVal* o, const char* key, float* dst, int n, const float* dflt) {
        const JVal* v = o ? o->get(key) : nullptr;
        for (int i = 0; i < n; i++) dst[i] = dflt[i];
        if (v && v->kind == JVal::Arr && (int)v->arr.size() >= n)
            for (int i = 0; i < n; i++) if (v->arr[i].is_num()) dst[i] = (float)v->arr[i].num;
    }

    // ---- buffers: GLB BIN chunk (buffer 0 without uri), data: URIs, files relative to base_dir
    bool load_buffers(const std::vector<uint8_t>* glb_bin) {
        const JVal* bs = arr("buffers");
        if (!bs) return true;
        for (size_t i = 0; i < bs->arr.size(); i++) {
            const JVal& b = bs->arr[i];
            std::vector<uint8_t> data;
            const JVal* uri = b.get("uri");
            if (!uri || uri->kind != JVal::Str) {
                if (i == 0 && glb_bin) data = *glb_bin;
                else return fail("buffer " + std::to_string(i) + " has no uri and there is no GLB BIN chunk");
            } else if (uri->str.compare(0, 5, "data:") == 0) {
                size_t comma = uri->str.find(',');
                if (comma == std::string::npos || uri->str.find(";base64") == std::string::npos || uri->str.find(";base64") > comma)
                    return fail("buffer " + std::to_string(i) + ": only base64 data URIs are supported");
                if (!base64_decode(uri->str.c_str() + comma + 1, uri->str.size() - comma - 1, data))
                    return fail("buffer " + std::to_string(i) + ": invalid base64");
            } else {
                std::string path = (uri->str.size() && uri->str[0] == '/') ? uri->str : base_dir + "/" + uri->str;
                if (!read_file(path, data)) return fail("cannot read buffer file " + path);
            }
            size_t want = 0;
            if (get_index(&b, "byteLength", &want) && data.size() < want)
                return fail("buffer " + std::to_string(i) + " is shorter than its byteLength");
            buffers.push_back(std::move(data));
        }
        return true;
    }

    struct View { const std::vector<uint8_t>* data; size_t start, stride, count, elem_size; int component; };
    // gltf_loader.rs:497-507: start = view.offset + accessor.offset; stride = view.stride or accessor.size
    bool accessor_view(size_t acc_index, View* v) {
        const JVal* accs = arr("accessors");
        if (!accs || acc_index >= accs->arr.size()) return fail("accessor index out of range");
        const JVal& a = accs->arr[acc_index];
        size_t bv_index;
        if (!get_index(&a, "bufferView", &bv_index)) return fail("Accessor missing buffer view");
        const JVal* bvs = arr("bufferViews");
        if (!bvs || bv_index >= bvs->arr.size()) return fail("bufferView index out of range");
        const JVal& bv = bvs->arr[bv_index];
        size_t buf_index = 0, bv_off = 0, a_off = 0, count = 0, stride = 0, ctype = 0;
        if (!get_index(&bv, "buffer", &buf_index) || buf_index >= buffers.size()) return fail("buffer index out of range");
        get_index(&bv, "byteOffset", &bv_off);
        get_index(&a, "byteOffset", &a_off);
        if (!get_index(&a, "count", &count)) return fail("accessor without count");
        if (!get_index(&a, "componentType", &ctype)) return fail("accessor without componentType");
        const JVal* type = a.get("type");
        std::string t = (type && type->kind == JVal::Str) ? type->str : "SCALAR";
        size_t ncomp = t == "SCALAR" ? 1 : t == "VEC2" ? 2 : t == "VEC3" ? 3 : t == "VEC4" ? 4 : t == "MAT2" ? 4 : t == "MAT3" ? 9 : t == "MAT4" ? 16 : 0;
        size_t csize = (ctype == 5120 || ctype == 5121) ? 1 : (ctype == 5122 || ctype == 5123) ? 2 : (ctype == 5125 || ctype == 5126) ? 4 : 0;
        if (!ncomp || !csize) return fail("unsupported accessor type");
        v->elem_size = ncomp * csize;
        v->stride = get_index(&bv, "byteStride", &stride) && stride ? stride : v->elem_size;
        v->data = &buffers[buf_index]; v->start = bv_off + a_off; v->count = count; v->component = (int)ctype;
        return true;
    }
    // get_accessor_data::<[f32;3]>, gltf_loader.rs:497-542 (little-endian f32 triples; out-of-bounds is an error)
    bool read_positions(size_t acc_index, std::vector<float>& out3) {
        View v;
        if (!accessor_view(acc_index, &v)) return false;
        // the whole range is checked before anything is allocated (count, offsets and stride come from the file)
        const size_t size = v.data->size();
        if (v.elem_size < 12) return fail("Buffer access out of bounds");
        if (v.count && (v.start > size || v.elem_size > size - v.start || (v.count - 1) > (size - v.start - v.elem_size) / v.stride))
            return fail("Buffer access out of bounds");
        out3.resize(v.count * 3);
        for (size_t i = 0; i < v.count; i++) memcpy(&out3[i * 3], v.data->data() + v.start + i * v.stride, 12);
        return true;
    }
    // get_indices_data, gltf_loader.rs:545-590: tightly packed u8/u16/u32; entries past the buffer end are skipped
    bool read_indices(size_t acc_index, std::vector<uint32_t>& out) {
        View v;
        if (!accessor_view(acc_index, &v)) return false;
        size_t w = v.component == 5121 ? 1 : v.component == 5123 ? 2 : v.component == 5125 ? 4 : 0;
        if (!w) return fail("Unsupported index data type");
        out.clear();
        // packed entries: the ones past the buffer end form a tail, so "skipped" = the count is clipped
        const size_t size = v.data->size();
        size_t n_ok = (v.start <= size && w <= size - v.start) ? (size - v.start - w) / w + 1 : 0;
        if (n_ok > v.count) n_ok = v.count;
        out.reserve(n_ok);
        for (size_t i = 0; i < n_ok; i++) {
            uint32_t x = 0;
            memcpy(&x, v.data->data() + v.start + i * w, w);
            out.push_back(x);
        }
        return true;
    }

    // ---- textures: dimensions from the PNG / JPEG header, RGBA8 block zero-filled (not decoded)
    bool image_bytes(const JVal& img, std::vector<uint8_t>& bytes) {
        const JVal* uri = img.get("uri");
        size_t bv_index;
        if (uri && uri->kind == JVal::Str) {
            if (uri->str.compare(0, 5, "data:") == 0) {
                size_t comma = uri->str.find(',');
                return comma != std::string::npos && base64_decode(uri->str.c_str() + comma + 1, uri->str.size() - comma - 1, bytes);
            }
            return read_file(base_dir + "/" + uri->str, bytes);
        }
        if (get_index(&img, "bufferView", &bv_index)) {
            const JVal* bvs = arr("bufferViews");
            if (!bvs || bv_index >= bvs->arr.size()) return false;
            size_t b = 0, off = 0, len = 0;
            if (!get_index(&bvs->arr[bv_index], "buffer", &b) || b >= buffers.size()) return false;
            get_index(&bvs->arr[bv_index], "byteOffset", &off);
            get_index(&bvs->arr[bv_index], "byteLength", &len);
            if (off + len > buffers[b].size()) return false;
            bytes.assign(buffers[b].begin() + off, buffers[b].begin() + off + len);
            return true;
        }
        return false;
    }
    static bool image_size(const std::vector<uint8_t>& d, uint32_t* w, uint32_t* h) {
        if (d.size() >= 24 && !memcmp(d.data(), "\x89PNG\r\n\x1a\n", 8)) {
            *w = (uint32_t)d[16] << 24 | d[17] << 16 | d[18] << 8 | d[19];
            *h = (uint32_t)d[20] << 24 | d[21] << 16 | d[22] << 8 | d[23];
            return true;
        }
        if (d.size() >= 4 && d[0] == 0xFF && d[1] == 0xD8) {
            size_t i = 2;
            while (i + 9 < d.size()) {
                if (d[i] != 0xFF) { i++; continue; }
                uint8_t m = d[i + 1];
                if (m >= 0xC0 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
                    *h = (uint32_t)d[i + 5] << 8 | d[i + 6]; *w = (uint32_t)d[i + 7] << 8 | d[i + 8];
                    return true;
                }
                i += 2 + ((size_t)d[i + 2] << 8 | d[i + 3]);
            }
        }
        return false;
    }
    void process_textures() { // gltf_loader.rs:128-184
        const JVal* texs = arr("textures");
        const JVal* imgs = arr("images");
        if (!texs) return;
        for (auto& t : texs->arr) {
            size_t src;
            if (!get_index(&t, "source", &src) || !imgs || src >= imgs->arr.size()) continue;
            std::vector<uint8_t> bytes;
            uint32_t w = 0, h = 0;
            if (!image_bytes(imgs->arr[src], bytes) || !image_size(bytes, &w, &h)) {
                out->warnings.push_back("image " + std::to_string(src) + ": payload unreadable, texture recorded as 0x0");
                w = h = 0;
            }
            if ((uint64_t)w * h * 4u > (1ull << 30) || out->texture_data.size() + (uint64_t)w * h * 4u > 0xFFFFFFFFull) {
                // offsets and sizes are u32 (TextureInfo); a header claiming more is not given a zero-filled block
                out->warnings.push_back("image " + std::to_string(src) + ": " + std::to_string(w) + "x" + std::to_string(h) + " exceeds the texture block limits, texture recorded as 0x0");
                w = h = 0;
            }
            RtTextureInfo ti;
            memset(&ti, 0, sizeof ti);
            ti.width = w; ti.height = h; ti.format = 3; ti.mip_levels = 1; // TextureInfo::new(w, h, 3 /*RGBA8*/, offset, size)
            ti.offset = (uint32_t)out->texture_data.size();
            ti.size = w * h * 4u;
            out->textures.push_back(ti);
            out->texture_data.resize(out->texture_data.size() + (size_t)ti.size, 0);
        }
    }

    // ---- materials, gltf_loader.rs:397-489
    void convert_material(const JVal& m, RtMaterial* mat) {
        const JVal* ext = m.get("extensions");
        const JVal* sg = ext ? ext->get("KHR_materials_pbrSpecularGlossiness") : nullptr;
        const JVal* pbr = m.get("pbrMetallicRoughness");
        const float one4[4] = {1, 1, 1, 1}, one3[3] = {1, 1, 1}, zero3[3] = {0, 0, 0};
        if (sg) { // Material::specular_glossiness, shared/src/lib.rs:335-346
            float diffuse[4], specular[3];
            vec_or(sg, "diffuseFactor", diffuse, 4, one4);
            vec_or(sg, "specularFactor", specular, 3, one3);
            float gloss = num_or(sg, "glossinessFactor", 1.0f);
            rt_material_new(mat, diffuse, 0.0f, 1.0f - gloss, zero3, 1.5f, 0.0f);
            mat->material_type = 1;
            memcpy(mat->diffuse_factor, diffuse, 12);
            memcpy(mat->specular_color, specular, 12);
            mat->glossiness_factor = gloss;
        } else {
            float base[4];
            vec_or(pbr, "baseColorFactor", base, 4, one4);
            rt_material_new(mat, base, num_or(pbr, "metallicFactor", 1.0f), num_or(pbr, "roughnessFactor", 1.0f), zero3, 1.5f, 0.0f);
        }
        vec_or(&m, "emissiveFactor", mat->emission, 3, zero3);
        if (const JVal* tr = ext ? ext->get("KHR_materials_transmission") : nullptr) { // set_transmission, :395-399
            uint16_t h = rt_f32_to_f16(num_or(tr, "transmissionFactor", 0.0f));
            mat->ior_transmission_f16 = (mat->ior_transmission_f16 & 0x0000FFFFu) | ((uint32_t)h << 16);
        }
        if (const JVal* io = ext ? ext->get("KHR_materials_ior") : nullptr) { // set_ior, :387-391
            uint16_t h = rt_f32_to_f16(num_or(io, "ior", 1.5f));
            mat->ior_transmission_f16 = (mat->ior_transmission_f16 & 0xFFFF0000u) | h;
        }
        if (const JVal* sp = ext ? ext->get("KHR_materials_specular") : nullptr) {
            mat->specular_factor = num_or(sp, "specularFactor", 1.0f);
            vec_or(sp, "specularColorFactor", mat->specular_color, 3, one3);
        }
        if (const JVal* vo = ext ? ext->get("KHR_materials_volume") : nullptr) {
            mat->thickness_factor = num_or(vo, "thicknessFactor", 0.0f);
            mat->attenuation_distance = num_or(vo, "attenuationDistance", INFINITY);
            vec_or(vo, "attenuationColor", mat->attenuation_color, 3, one3);
        }
        uint32_t ti[8];
        for (int i = 0; i < 8; i++) ti[i] = 0xFFFFFFFFu;
        int n = 0;
        auto add = [&](const JVal* holder, const char* key) {
            const JVal* t = holder ? holder->get(key) : nullptr;
            size_t idx;
            if (t && get_index(t, "index", &idx) && n < 8) ti[n++] = (uint32_t)idx;
        };
        add(pbr, "baseColorTexture");
        add(pbr, "metallicRoughnessTexture");
        add(&m, "normalTexture");
        add(&m, "emissiveTexture");
        memcpy(mat->texture_indices, ti, sizeof ti);
    }

    // ---- nodes
    M4 node_matrix(const JVal& node) {
        const JVal* mx = node.get("matrix");
        if (mx && mx->kind == JVal::Arr && mx->arr.size() == 16) {
            M4 m;
            for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) m.c[c][r] = (float)mx->arr[c * 4 + r].num;
            return m;
        }
        const float t0[3] = {0, 0, 0}, r0[4] = {0, 0, 0, 1}, s0[3] = {1, 1, 1};
        float t[3], q[4], s[3];
        vec_or(&node, "translation", t, 3, t0);
        vec_or(&node, "rotation", q, 4, r0);
        vec_or(&node, "scale", s, 3, s0);
        M4 T = m4_identity(), R = m4_identity(), S = m4_identity();
        T.c[3][0] = t[0]; T.c[3][1] = t[1]; T.c[3][2] = t[2];
        S.c[0][0] = s[0]; S.c[1][1] = s[1]; S.c[2][2] = s[2];
        { // quaternion (x, y, z, w) -> rotation matrix (cgmath-style doubling)
            float x = q[0], y = q[1], z = q[2], w = q[3];
            float x2 = x + x, y2 = y + y, z2 = z + z;
            float xx2 = x2 * x, xy2 = x2 * y, xz2 = x2 * z, yy2 = y2 * y, yz2 = y2 * z, zz2 = z2 * z;
            float sy2 = y2 * w, sz2 = z2 * w, sx2 = x2 * w;
            R.c[0][0] = 1.0f - yy2 - zz2; R.c[0][1] = xy2 + sz2; R.c[0][2] = xz2 - sy2;
            R.c[1][0] = xy2 - sz2; R.c[1][1] = 1.0f - xx2 - zz2; R.c[1][2] = yz2 + sx2;
            R.c[2][0] = xz2 + sy2; R.c[2][1] = yz2 - sx2; R.c[2][2] = 1.0f - xx2 - yy2;
        }
        return m4_mul(m4_mul(T, R), S);
    }

    bool process_primitive(const JVal& prim, const M4& xf) { // gltf_loader.rs:290-394
        size_t mi;
        uint32_t material_id = 0;
        if (get_index(&prim, "material", &mi)) { auto it = material_map.find(mi); if (it != material_map.end()) material_id = it->second; }
        const JVal* attrs = prim.get("attributes");
        size_t pos_acc;
        if (!get_index(attrs, "POSITION", &pos_acc)) return fail("Primitive missing position data");
        std::vector<float> pos;
        if (!read_positions(pos_acc, pos)) return false;
        const size_t npos = pos.size() / 3;
        std::unordered_map<PosKey, uint32_t, PosHash> vertex_map; // per primitive, like the reference
        bool oob = false;
        auto vertex_index = [&](size_t i) -> uint32_t {
            if (i >= npos) { oob = true; return 0; } // the reference panics on this index
            float tp[3];
            m4_point(xf, &pos[i * 3], tp);
            PosKey k;
            memcpy(k.b, tp, 12);
            auto it = vertex_map.find(k);
            if (it != vertex_map.end()) return it->second;
            uint32_t ni = (uint32_t)out->vertices.size();
            RtVertex v; memcpy(v.position, tp, 12);
            out->vertices.push_back(v);
            vertex_map.emplace(k, ni);
            return ni;
        };
        auto tri = [&](uint32_t a, uint32_t b, uint32_t c) { RtTriangle t; t.v0_index = a; t.v1_index = b; t.v2_index = c; t.material_id = material_id; out->triangles.push_back(t); };
        size_t mode = 4, idx_acc;
        get_index(&prim, "mode", &mode);
        if (mode == 4) { // TRIANGLES
            if (get_index(&prim, "indices", &idx_acc)) {
                std::vector<uint32_t> idx;
                if (!read_indices(idx_acc, idx)) return false;
                for (size_t i = 0; i + 2 < idx.size(); i += 3) {
                    uint32_t a = vertex_index(idx[i]), b = vertex_index(idx[i + 1]), c = vertex_index(idx[i + 2]);
                    tri(a, b, c);
                }
            } else {
                for (size_t i = 0; i + 2 < npos; i += 3) {
                    uint32_t a = vertex_index(i), b = vertex_index(i + 1), c = vertex_index(i + 2);
                    tri(a, b, c);
                }
            }
        } else if (mode == 6) { // TRIANGLE_FAN over the position array (indices ignored, as in the reference)
            if (npos >= 3) {
                uint32_t c0 = vertex_index(0);
                for (size_t i = 1; i + 1 < npos; i++) { uint32_t a = vertex_index(i), b = vertex_index(i + 1); tri(c0, a, b); }
            }
        } else if (mode == 5) { // TRIANGLE_STRIP, alternating winding
            for (size_t i = 0; i + 2 < npos; i++) {
                uint32_t a = vertex_index(i), b = vertex_index(i + 1), c = vertex_index(i + 2);
                if (i % 2 == 0) tri(a, b, c); else tri(a, c, b);
            }
        } else {
            out->warnings.push_back("Unsupported primitive mode: " + std::to_string(mode));
        }
        if (oob) return fail("index refers to a vertex beyond the POSITION accessor");
        return true;
    }

    bool process_node(size_t node_index, const M4& parent, int depth) { // gltf_loader.rs:188-226
        const JVal* nodes = arr("nodes");
        if (!nodes || node_index >= nodes->arr.size()) return fail("node index out of range");
        if (depth > 256) return fail("node hierarchy deeper than 256 (cycle?)");
        const JVal& node = nodes->arr[node_index];
        M4 xf = m4_mul(parent, node_matrix(node));
        size_t mesh_i, cam_i;
        if (get_index(&node, "mesh", &mesh_i)) {
            const JVal* meshes = arr("meshes");
            if (!meshes || mesh_i >= meshes->arr.size()) return fail("mesh index out of range");
            const JVal* prims = meshes->arr[mesh_i].get("primitives");
            if (prims && prims->kind == JVal::Arr)
                for (auto& pr : prims->arr) if (!process_primitive(pr, xf)) return false;
        }
        const float origin[3] = {0, 0, 0}, neg_z[3] = {0, 0, -1}, pos_y[3] = {0, 1, 0};
        if (get_index(&node, "camera", &cam_i)) { // convert_camera, :228-251
            const JVal* cams = arr("cameras");
            if (!cams || cam_i >= cams->arr.size()) return fail("camera index out of range");
            RtCamera cam;
            m4_point(xf, origin, cam.position);
            m4_vector(xf, neg_z, cam.direction); normalize(cam.direction);
            m4_vector(xf, pos_y, cam.up); normalize(cam.up);
            cam.fov = 45.0f;
            const JVal* persp = cams->arr[cam_i].get("perspective");
            const JVal* type = cams->arr[cam_i].get("type");
            if (persp && !(type && type->kind == JVal::Str && type->str == "orthographic")) {
                float yfov = num_or(persp, "yfov", 0.7853982f);
                cam.fov = yfov * (180.0f / 3.14159265358979323846f); // f32::to_degrees
            }
            out->cameras.push_back(cam);
        }
        const JVal* next = node.get("extensions");
        const JVal* lp = next ? next->get("KHR_lights_punctual") : nullptr;
        size_t light_i;
        if (lp && get_index(lp, "light", &light_i)) { // convert_light, :254-287
            const JVal* dext = doc.get("extensions");
            const JVal* dl = dext ? dext->get("KHR_lights_punctual") : nullptr;
            const JVal* ls = dl ? dl->get("lights") : nullptr;
            if (!ls || ls->kind != JVal::Arr || light_i >= ls->arr.size()) return fail("light index out of range");
            const JVal& l = ls->arr[light_i];
            float position[3], direction[3], color[3];
            const float one3[3] = {1, 1, 1};
            m4_point(xf, origin, position);
            m4_vector(xf, neg_z, direction); normalize(direction);
            vec_or(&l, "color", color, 3, one3);
            float intensity = num_or(&l, "intensity", 1.0f);
            float range = num_or(&l, "range", INFINITY);
            const JVal* type = l.get("type");
            std::string t = (type && type->kind == JVal::Str) ? type->str : "point";
            RtLight light;
            if (t == "directional") rt_light_directional(&light, direction, color, intensity);
            else if (t == "spot") {
                const JVal* sp = l.get("spot");
                rt_light_spot(&light, position, direction, color, intensity, range, num_or(sp, "innerConeAngle", 0.0f),
                              num_or(sp, "outerConeAngle", 0.7853982f));
            } else rt_light_point(&light, position, color, intensity, range);
            out->lights.push_back(light);
        }
        const JVal* ch = node.get("children");
        if (ch && ch->kind == JVal::Arr)
            for (auto& c : ch->arr) if (c.is_num() && !process_node((size_t)c.num, xf, depth + 1)) return false;
        return true;
    }

    bool extract(int scene_index) { // extract_scene, gltf_loader.rs:77-125
        const JVal* scenes = arr("scenes");
        out->scene_count = scenes ? (uint32_t)scenes->arr.size() : 0;
        size_t si = 0;
        if (scene_index >= 0) {
            if (!scenes || (size_t)scene_index >= scenes->arr.size()) return fail("Scene " + std::to_string(scene_index) + " not found");
            si = (size_t)scene_index;
        } else {
            if (!scenes || scenes->arr.empty()) return fail("No scenes found in glTF file");
            size_t d;
            if (get_index(&doc, "scene", &d) && d < scenes->arr.size()) si = d;
        }
        process_textures();
        if (const JVal* mats = arr("materials"))
            for (size_t i = 0; i < mats->arr.size(); i++) {
                RtMaterial m;
                convert_material(mats->arr[i], &m);
                material_map[i] = (uint32_t)out->materials.size();
                out->materials.push_back(m);
            }
        const JVal* roots = scenes->arr[si].get("nodes");
        M4 id = m4_identity();
        if (roots && roots->kind == JVal::Arr)
            for (auto& r : roots->arr) if (r.is_num() && !process_node((size_t)r.num, id, 0)) return false;
        return true;
    }
};

int load_impl(const uint8_t* data, uint64_t n, const char* base_dir, int scene_index, RtGltfScene** out, char* err, size_t err_len) {
    auto set_err = [&](const std::string& m) { if (err && err_len) { snprintf(err, err_len, "%s", m.c_str()); } return -1; };
    if (!data || !out) return set_err("NULL argument");
    Loader L;
    L.base_dir = (base_dir && *base_dir) ? base_dir : ".";
    std::vector<uint8_t> bin;
    bool have_bin = false;
    const char* json = (const char*)data;
    size_t json_len = (size_t)n;
    if (n >= 12 && !memcmp(data, "glTF", 4)) { // GLB container: header, JSON chunk, optional BIN chunk
        uint32_t version, total;
        memcpy(&version, data + 4, 4); memcpy(&total, data + 8, 4);
        if (version != 2) return set_err("unsupported GLB version");
        if (total > n) return set_err("GLB length field exceeds the data");
        size_t off = 12;
        bool have_json = false;
        while (off + 8 <= total) {
            uint32_t clen, ctype;
            memcpy(&clen, data + off, 4); memcpy(&ctype, data + off + 4, 4);
            off += 8;
            if (off + clen > total) return set_err("GLB chunk exceeds the file");
            if (ctype == 0x4E4F534Au && !have_json) { json = (const char*)data + off; json_len = clen; have_json = true; }
            else if (ctype == 0x004E4942u && !have_bin) { bin.assign(data + off, data + off + clen); have_bin = true; }
            off += (clen + 3u) & ~3u;
        }
        if (!have_json) return set_err("GLB without a JSON chunk");
    }
    JParser P{json, json + json_len, {}};
    if (!P.parse(L.doc) || L.doc.kind != JVal::Obj) return set_err("glTF JSON: " + (P.err.empty() ? std::string("not an object") : P.err));
    std::unique_ptr<RtGltfScene> scene(new RtGltfScene());
    L.out = scene.get();
    if (!L.load_buffers(have_bin ? &bin : nullptr) || !L.extract(scene_index)) return set_err(L.err);
    *out = scene.release();
    return 0;
}

} // namespace

extern "C" {

// no exception crosses the C ABI: allocation failures and anything else thrown become an error string
static int load_guarded(const uint8_t* data, uint64_t n, const char* base_dir, int scene_index, RtGltfScene** out, char* err, size_t err_len) {
    try {
        return load_impl(data, n, base_dir, scene_index, out, err, err_len);
    } catch (const std::exception& e) {
        if (err && err_len) snprintf(err, err_len, "glTF load failed: %s", e.what());
    } catch (...) {
        if (err && err_len) snprintf(err, err_len, "glTF load failed");
    }
    return -1;
}

int rt_gltf_load_memory(const uint8_t* data, uint64_t n, const char* base_dir, int scene_index, RtGltfScene** out, char* err, uint64_t err_len) {
    return load_guarded(data, n, base_dir, scene_index, out, err, (size_t)err_len);
}

int rt_gltf_load_file(const char* path, int scene_index, RtGltfScene** out, char* err, uint64_t err_len) {
    std::vector<uint8_t> data;
    if (!path || !read_file(path, data)) {
        if (err && err_len) snprintf(err, (size_t)err_len, "cannot read %s", path ? path : "(null)");
        return -1;
    }
    std::string p(path);
    size_t slash = p.find_last_of('/');
    std::string dir = slash == std::string::npos ? "." : p.substr(0, slash);
    return load_guarded(data.data(), data.size(), dir.c_str(), scene_index, out, err, (size_t)err_len);
}

void rt_gltf_free(RtGltfScene* s) { delete s; }

void rt_gltf_counts(const RtGltfScene* s, RtGltfCounts* c) {
    if (!s || !c) return;
    c->triangles = s->triangles.size(); c->vertices = s->vertices.size(); c->materials = s->materials.size();
    c->spheres = s->spheres.size(); c->lights = s->lights.size(); c->cameras = s->cameras.size();
    c->textures = s->textures.size(); c->texture_bytes = s->texture_data.size(); c->scenes = s->scene_count;
    c->warnings = s->warnings.size();
}
const RtTriangle* rt_gltf_triangles(const RtGltfScene* s) { return s ? s->triangles.data() : nullptr; }
const RtVertex* rt_gltf_vertices(const RtGltfScene* s) { return s ? s->vertices.data() : nullptr; }
const RtMaterial* rt_gltf_materials(const RtGltfScene* s) { return s ? s->materials.data() : nullptr; }
const RtSphere* rt_gltf_spheres(const RtGltfScene* s) { return s ? s->spheres.data() : nullptr; }
const RtLight* rt_gltf_lights(const RtGltfScene* s) { return s ? s->lights.data() : nullptr; }
const RtCamera* rt_gltf_cameras(const RtGltfScene* s) { return s ? s->cameras.data() : nullptr; }
const RtTextureInfo* rt_gltf_textures(const RtGltfScene* s) { return s ? s->textures.data() : nullptr; }
const uint8_t* rt_gltf_texture_data(const RtGltfScene* s) { return s ? s->texture_data.data() : nullptr; }
const char* rt_gltf_warning(const RtGltfScene* s, uint64_t i) { return (s && i < s->warnings.size()) ? s->warnings[i].c_str() : nullptr; }

} // extern "C"
