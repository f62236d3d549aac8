// hj_gltf.h — render_option.json and glTF 2.0 ingest of the headless loader, with the semantics of the reference's
// loaders: /root/reference/include/loader/render_json_loader.h:97-202 (+ fps.txt, :16-60) and
// include/loader/gltfloader.h:1068-1601 — one instance per mesh node, de-indexed triangles with an identity index
// buffer (:1484-1491), geometric normals / zero uvs when absent (:1460-1482), the KHR clearcoat / sheen / transmission /
// ior / emissive_strength and custom "ThinFilm" material extensions (:1194-1258), ideal_specular and is_light as the
// loader derives them (:1162-1167, :1260-1263), key-0 injection of every node's static TRS (:1313-1343), the camera node
// (:1516-1524) and the animation samplers indexed by CHANNEL index (sic, :1541-1542).
#pragma once
#include <fstream>
#include <map>
#include <sstream>
#include "hj_host_scene.h"
#include "hj_json.h"
#include "hj_png.h"

namespace hj {

inline std::string read_text(const std::string& path) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open " + path);
    std::stringstream ss; ss << f.rdbuf();
    return ss.str();
}

// render_json_loader.h:97-202. `dir` = directory the relative paths in the file are resolved against (the reference
// resolves them against the working directory, which is the directory of render_option.json in its layout).
inline RenderOption load_render_option(const std::string& path, const std::string& fps_path) {
    const Json j = JsonParser::parse(read_text(path));
    RenderOption o;
    const Json& im = j.at("Image");
    o.image_width = (int)im.at("image_width").number(); o.image_height = (int)im.at("image_height").number();
    o.image_name = im.at("image_name").string(); o.image_directory = im.at("image_directory").string();
    o.max_spp = (unsigned)im.at("max_spp").number();
    o.gltf_path = j.at("GLTF_file").at("gltf_filepath").string(); o.gltf_name = j.at("GLTF_file").at("gltf_filename").string();
    o.render_mode = j.at("Render_mode").string();
    const Json& cam = j.at("Camera");
    for (int k = 0; k < 3; k++) { o.camera_position[k] = (float)cam.at("camera_position").at(k).number(); o.camera_direction[k] = (float)cam.at("camera_direction").at(k).number(); }
    o.camera_fov = (float)(M_PI * cam.at("camera_fov").number() / 180.0f);   // :146, double maths, stored as float
    o.allow_camera_animation = cam.at("allow_camera_animation").boolean();
    o.ptxfile_path = j.at("PTX_File").at("ptxfile_path").string();     // kept and ignored: there is no device module to load
    const Json& an = j.at("Animation");
    o.fps = (unsigned)an.at("fps").number(); o.start_frame = (int)an.at("start_frame").number(); o.end_frame = (int)an.at("end_frame").number();
    o.time_limit = (float)an.at("time_limit").number();
    {   // fpsLoader (:16-60): a single unsigned in ./fps.txt overrides Animation.fps
        std::ifstream f(fps_path);
        unsigned v = 0;
        if (f && (f >> v) && v > 0) o.fps = v;
    }
    const Json& sky = j.at("Sky");
    o.IBL_path = sky.at("IBL_path").string(); o.IBL_intensity = (float)sky.at("IBL_intensity").number(); o.use_IBL = sky.at("use_IBL").boolean();
    for (int k = 0; k < 3; k++) o.scene_sky_default[k] = (float)sky.at("scene_sky_default").at(k).number();
    o.use_date = j.at("Option").at("use_date").boolean(); o.save_renderOption = j.at("Option").at("save_renderOption").boolean();
    o.LUT_path = j.at("LUT").at("LUT_path").string();
    if (const Json* ig = j.find("Integrator")) {                       // extension of this loader, defaults = reference behaviour
        const std::string m = ig->string_or("mode", "MIS");
        o.integrator = m == "PT" ? HJ_INTEGRATOR_PT : (m == "NEE" ? HJ_INTEGRATOR_NEE : HJ_INTEGRATOR_MIS);
        o.max_depth = ig->int_or("max_depth", 10);
        const std::string t = ig->string_or("tonemap", "none");
        o.tonemap = t == "ACES" ? HJ_TONEMAP_ACES : (t == "Uchimura" ? HJ_TONEMAP_UCHIMURA : HJ_TONEMAP_NONE);
    }
    return o;
}

namespace detail {
inline std::vector<uint8_t> base64_decode(const std::string& s, size_t from) {
    std::vector<uint8_t> out; uint32_t acc = 0; int bits = 0;
    for (size_t i = from; i < s.size(); i++) {
        const char c = s[i]; int v;
        if (c >= 'A' && c <= 'Z') v = c - 'A'; else if (c >= 'a' && c <= 'z') v = c - 'a' + 26; else if (c >= '0' && c <= '9') v = c - '0' + 52;
        else if (c == '+') v = 62; else if (c == '/') v = 63; else continue;
        acc = (acc << 6) | (uint32_t)v; bits += 6;
        if (bits >= 8) { bits -= 8; out.push_back((uint8_t)(acc >> bits)); }
    }
    return out;
}
struct Accessor { const uint8_t* ptr; size_t count, stride; int comp_type, ncomp; };
inline int comp_size(int t) { return (t == 5120 || t == 5121) ? 1 : (t == 5122 || t == 5123) ? 2 : 4; }
inline Accessor accessor(const Json& g, const std::vector<std::vector<uint8_t>>& buffers, int idx) {
    const Json& a = g.at("accessors").at((size_t)idx);
    const Json& bv = g.at("bufferViews").at((size_t)a.at("bufferView").number());
    Accessor r;
    r.comp_type = (int)a.at("componentType").number();
    const std::string& ty = a.at("type").string();
    r.ncomp = ty == "SCALAR" ? 1 : ty == "VEC2" ? 2 : ty == "VEC3" ? 3 : ty == "VEC4" ? 4 : ty == "MAT4" ? 16 : 0;
    if (!r.ncomp) throw std::runtime_error("gltf: unsupported accessor type " + ty);
    r.count = (size_t)a.at("count").number();
    const size_t off = (size_t)bv.number_or("byteOffset", 0) + (size_t)a.number_or("byteOffset", 0);
    const size_t bs = (size_t)bv.number_or("byteStride", 0);
    r.stride = bs ? bs : (size_t)comp_size(r.comp_type) * r.ncomp;       // Accessor::ByteStride
    const auto& buf = buffers.at((size_t)bv.at("buffer").number());
    if (r.count && off + (r.count - 1) * r.stride + (size_t)comp_size(r.comp_type) * r.ncomp > buf.size()) throw std::runtime_error("gltf: accessor out of bounds");
    r.ptr = buf.data() + off;
    return r;
}
// the loader reads POSITION / NORMAL / TEXCOORD_0 / sampler data as 32-bit floats (tinygltf hands the reference raw bytes that it casts to
// float*, gltfloader.h:1425-1482,1555-1590): anything else would be read past the range accessor() validated
inline void require_float(const Accessor& a, int ncomp, const char* what) {
    if (a.comp_type != 5126 || a.ncomp != ncomp) throw std::runtime_error(std::string("gltf: ") + what + " must be FLOAT with " + std::to_string(ncomp) + " component(s)");
}
inline float get_f(const Accessor& a, size_t i, int c) { float v; memcpy(&v, a.ptr + i * a.stride + 4 * (size_t)c, 4); return v; }
inline uint32_t get_index(const Accessor& a, size_t i) {
    const uint8_t* p = a.ptr + i * a.stride;
    switch (a.comp_type) {
        case 5120: return (uint32_t)(int)*(const int8_t*)p;
        case 5121: return *p;
        case 5122: { int16_t v; memcpy(&v, p, 2); return (uint32_t)(int)v; }
        case 5123: { uint16_t v; memcpy(&v, p, 2); return v; }
        default: { uint32_t v; memcpy(&v, p, 4); return v; }
    }
}
}  // namespace detail

// gltfloader.h:1068-1601. Fills `sc`; may overwrite the camera fields of `opt` (camera node, :1516-1524).
inline void load_gltf(const std::string& dir, const std::string& name, RenderOption& opt, HostScene& sc) {
    using namespace detail;
    const Json g = JsonParser::parse(read_text(dir + name));
    std::vector<std::vector<uint8_t>> buffers;
    if (const Json* bs = g.find("buffers"))
        for (const Json& b : bs->arr) {
            const std::string uri = b.string_or("uri", "");
            if (uri.compare(0, 5, "data:") == 0) { const size_t c = uri.find(','); buffers.push_back(base64_decode(uri, c == std::string::npos ? 0 : c + 1)); }
            else buffers.push_back(read_file(dir + uri));
        }
    // ---- materials (:1125-1266); textures are cached by uri (loader/texture_load.h:7-20)
    std::map<std::string, int> known_tex;
    auto load_texture = [&](int tex_index, int srgb) -> int {
        const Json& t = g.at("textures").at((size_t)tex_index);
        const std::string uri = g.at("images").at((size_t)t.at("source").number()).at("uri").string();
        auto it = known_tex.find(uri);
        if (it != known_tex.end()) return it->second;
        Image8 im = png_read(dir + uri);
        HostTexture ht; ht.width = im.width; ht.height = im.height; ht.srgb = srgb; ht.rgba = std::move(im.rgba); ht.name = uri;
        const int id = (int)sc.textures.size();
        sc.textures.push_back(std::move(ht));
        known_tex[uri] = id;
        return id;
    };
    if (const Json* mats = g.find("materials"))
        for (const Json& m : mats->arr) {
            hj_material o{};
            static const Json empty_obj = [] { Json e; e.type = Json::Object; return e; }();
            const Json& pbr = m.has("pbrMetallicRoughness") ? m.at("pbrMetallicRoughness") : empty_obj;
            for (int k = 0; k < 3; k++) o.basecolor[k] = pbr.has("baseColorFactor") ? (float)pbr.at("baseColorFactor").at((size_t)k).number() : 1.0f;
            o.basecolor_tex = pbr.has("baseColorTexture") ? load_texture((int)pbr.at("baseColorTexture").at("index").number(), 1) : -1;
            o.roughness = (float)pbr.number_or("roughnessFactor", 1.0);
            o.roughness_tex = pbr.has("metallicRoughnessTexture") ? load_texture((int)pbr.at("metallicRoughnessTexture").at("index").number(), 0) : -1;
            o.metallic = (float)pbr.number_or("metallicFactor", 1.0);
            o.metallic_tex = o.roughness_tex;                                                         // :1156
            for (int k = 0; k < 3; k++) o.emission[k] = m.has("emissiveFactor") ? (float)m.at("emissiveFactor").at((size_t)k).number() : 0.0f;
            o.emission_tex = -1;
            o.is_light = (o.emission[0] + o.emission[1] + o.emission[2] > 0.0) ? 1 : 0;            // :1162-1167, before the strength
            o.normal_tex = m.has("normalTexture") ? load_texture((int)m.at("normalTexture").at("index").number(), 0) : -1;
            o.sheen = 0; o.clearcoat = 0; o.clearcoat_tex = -1; o.specular[0] = o.specular[1] = o.specular[2] = 0;
            o.transmission = 0; o.bump_tex = -1; o.ior = 1.0; o.is_thinfilm = 0;
            if (const Json* ext = m.find("extensions")) {
                for (auto& kv : ext->obj) {
                    const Json& e = kv.second;
                    if (kv.first == "KHR_materials_clearcoat") { if (e.has("clearcoatFactor")) o.clearcoat = (float)e.at("clearcoatFactor").number(); }
                    else if (kv.first == "KHR_materials_sheen") { if (e.has("sheenRoughnessFactor")) o.sheen = (float)e.at("sheenRoughnessFactor").number(); }   // sic :1211
                    else if (kv.first == "KHR_materials_transmission") { if (e.has("transmissionFactor")) o.transmission = (float)e.at("transmissionFactor").number(); }
                    else if (kv.first == "KHR_materials_ior") { if (e.has("ior")) o.ior = (float)e.at("ior").number(); }
                    else if (kv.first == "KHR_materials_emissive_strength") {
                        if (e.has("emissiveStrength")) { const float s = (float)e.at("emissiveStrength").number(); for (int k = 0; k < 3; k++) o.emission[k] *= s; }   // float3 *= float
                    } else if (kv.first == "ThinFilm") { if (e.has("is_ThinFilm")) o.is_thinfilm = 1; }   // presence, not value (:1250-1256)
                }
            }
            o.ideal_specular = (o.roughness == 0 && o.transmission > 0) ? 1 : 0;                     // :1260-1263
            sc.materials.push_back(o);
        }
    // ---- nodes: key-0 TRS, meshes -> instances, camera (:1304-1528)
    const Json& nodes = g.at("nodes");
    sc.animations.assign(nodes.size(), NodeAnimation());
    for (size_t ni = 0; ni < nodes.size(); ni++) {
        const Json& n = nodes.at(ni);
        NodeAnimation& a = sc.animations[ni];
        a.t_key.push_back(0); a.r_key.push_back(0); a.s_key.push_back(0);
        const float dt[3] = {0, 0, 0}, dr[4] = {0, 0, 0, 1}, ds[3] = {1, 1, 1};
        for (int k = 0; k < 3; k++) a.t_val.push_back(n.has("translation") ? (float)n.at("translation").at((size_t)k).number() : dt[k]);
        for (int k = 0; k < 4; k++) a.r_val.push_back(n.has("rotation") ? (float)n.at("rotation").at((size_t)k).number() : dr[k]);
        for (int k = 0; k < 3; k++) a.s_val.push_back(n.has("scale") ? (float)n.at("scale").at((size_t)k).number() : ds[k]);
        if (n.has("mesh")) {
            {   // a mesh without a single triangle would give an empty [prim_offset, next) range, which hj_scene_upload refuses for the whole
                // scene ("strictly increasing"): such a node contributes no instance
                size_t tris = 0;
                for (const Json& prim : g.at("meshes").at((size_t)n.at("mesh").number()).at("primitives").arr)
                    tris += (size_t)g.at("accessors").at((size_t)prim.at("indices").number()).at("count").number() / 3;
                if (tris == 0) continue;
            }
            sc.prim_offsets.push_back((uint32_t)(sc.indices.size() / 3));
            sc.instance_node.push_back((uint32_t)ni);
            uint32_t prim_id = (uint32_t)(sc.vertices.size() / 9);
            const Json& mesh = g.at("meshes").at((size_t)n.at("mesh").number());
            for (const Json& prim : mesh.at("primitives").arr) {
                const Accessor ia = accessor(g, buffers, (int)prim.at("indices").number());
                const Json& at = prim.at("attributes");
                const Accessor pa = accessor(g, buffers, (int)at.at("POSITION").number());
                require_float(pa, 3, "POSITION");
                if (ia.ncomp != 1 || ia.comp_type == 5126) throw std::runtime_error("gltf: indices must be an integer SCALAR accessor");
                const bool has_n = at.has("NORMAL"), has_t = at.has("TEXCOORD_0");
                Accessor na{}, ta{};
                if (has_n) { na = accessor(g, buffers, (int)at.at("NORMAL").number()); require_float(na, 3, "NORMAL"); if (na.count < pa.count) throw std::runtime_error("gltf: NORMAL shorter than POSITION"); }
                if (has_t) { ta = accessor(g, buffers, (int)at.at("TEXCOORD_0").number()); require_float(ta, 2, "TEXCOORD_0"); if (ta.count < pa.count) throw std::runtime_error("gltf: TEXCOORD_0 shorter than POSITION"); }
                const uint32_t mat = (uint32_t)prim.int_or("material", 0);
                for (size_t t = 0; t < ia.count / 3; t++) {
                    uint32_t idx[3]; float v[3][3], nn[3][3], tc[3][2];
                    for (int k = 0; k < 3; k++) {
                        idx[k] = get_index(ia, 3 * t + (size_t)k);
                        if (idx[k] >= pa.count) throw std::runtime_error("gltf: vertex index out of range");
                        for (int c = 0; c < 3; c++) v[k][c] = get_f(pa, idx[k], c);
                    }
                    if (has_n) { for (int k = 0; k < 3; k++) for (int c = 0; c < 3; c++) nn[k][c] = get_f(na, idx[k], c); }
                    else {      // normalize(cross(v1 - v0, v2 - v0)), sutil: v * (1 / sqrtf(dot))   (:1460-1466)
                        const float e1[3] = {v[1][0] - v[0][0], v[1][1] - v[0][1], v[1][2] - v[0][2]}, e2[3] = {v[2][0] - v[0][0], v[2][1] - v[0][1], v[2][2] - v[0][2]};
                        const float cx = e1[1] * e2[2] - e1[2] * e2[1], cy = e1[2] * e2[0] - e1[0] * e2[2], cz = e1[0] * e2[1] - e1[1] * e2[0];
                        const float inv = 1.0f / sqrtf(cx * cx + cy * cy + cz * cz);
                        for (int k = 0; k < 3; k++) { nn[k][0] = cx * inv; nn[k][1] = cy * inv; nn[k][2] = cz * inv; }
                    }
                    for (int k = 0; k < 3; k++) { tc[k][0] = has_t ? get_f(ta, idx[k], 0) : 0.0f; tc[k][1] = has_t ? get_f(ta, idx[k], 1) : 0.0f; }
                    for (int k = 0; k < 3; k++) {
                        sc.vertices.insert(sc.vertices.end(), v[k], v[k] + 3);
                        sc.normals.insert(sc.normals.end(), nn[k], nn[k] + 3);
                        sc.texcoords.insert(sc.texcoords.end(), tc[k], tc[k] + 2);
                        sc.indices.push_back((uint32_t)sc.indices.size());                           // identity (:1491)
                    }
                    sc.material_ids.push_back(mat);
                    if (mat < sc.materials.size() && sc.materials[mat].is_light) {                   // :1496-1502
                        sc.light_prim_ids.push_back(prim_id);
                        sc.light_prim_emission.insert(sc.light_prim_emission.end(), sc.materials[mat].emission, sc.materials[mat].emission + 3);
                    }
                    prim_id++;
                }
            }
        } else if (n.has("camera") && opt.allow_camera_animation) {                                  // :1516-1524
            opt.camera_position[0] = opt.camera_position[1] = opt.camera_position[2] = 0;
            opt.camera_direction[0] = 0; opt.camera_direction[1] = 0; opt.camera_direction[2] = -1;
            opt.camera_animation_id = (int)ni;
            opt.camera_fov = (float)g.at("cameras").at((size_t)n.at("camera").number()).at("perspective").at("yfov").number();
        }
    }
    // ---- animations: sampler taken by CHANNEL index (:1541-1542)
    if (const Json* anims = g.find("animations"))
        for (const Json& anim : anims->arr) {
            const Json& ch = anim.at("channels"); const Json& sm = anim.at("samplers");
            for (size_t i = 0; i < ch.size(); i++) {
                const Json& smp = sm.at(i);
                const Accessor keys = accessor(g, buffers, (int)smp.at("input").number());
                const Accessor data = accessor(g, buffers, (int)smp.at("output").number());
                const Json& tgt = ch.at(i).at("target");
                NodeAnimation& a = sc.animations.at((size_t)tgt.at("node").number());
                const std::string& path = tgt.at("path").string();
                require_float(keys, 1, "animation sampler input");
                require_float(data, path == "rotation" ? 4 : 3, "animation sampler output");
                const size_t n_keys = keys.count < data.count ? keys.count : data.count;     // never past the shorter of the two
                for (size_t k = 0; k < n_keys; k++) {
                    const float key = get_f(keys, k, 0);
                    if (path == "translation") { for (int c = 0; c < 3; c++) a.t_val.push_back(get_f(data, k, c)); a.t_key.push_back(key); }
                    else if (path == "rotation") { for (int c = 0; c < 4; c++) a.r_val.push_back(get_f(data, k, c)); a.r_key.push_back(key); }
                    else if (path == "scale") { for (int c = 0; c < 3; c++) a.s_val.push_back(get_f(data, k, c)); a.s_key.push_back(key); }
                }
            }
        }
}

}  // namespace hj
