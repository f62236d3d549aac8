// henjou_b200_cli — headless host of the B200 path tracer: reads the reference's render_option.json (+ ./fps.txt) and its
// glTF scene, runs the reference's frame loop (Renderer::initializeAndRender, /root/reference/include/renderer/
// renderer.h:1053-1318) over the C ABI of libhenjou_b200.so instead of OptiX, and writes `<image_name>_<NNN>.png` into
// the working directory (renderer.h:1285-1302). No window, no GL. The denoise step (renderer.h:1256-1283) is hj_post:
// Render_mode "Default" is a pass-through as in the reference; the Denoise modes run the a-trous stand-in for the OptiX network.
//
//   henjou_b200_cli [render_option.json] [--device N] [--spp N] [--frames BEGIN END] [--width W --height H]
//                   [--out-dir DIR] [--dump-scene FILE] [--rebuild] [--quiet]
// Relative paths inside render_option.json are resolved against the directory of that file (the reference resolves them
// against its working directory, which is that directory in its layout).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <string>
#include <vector>
#include "hj_gltf.h"
#include "hj_hdr.h"

using namespace hj;

static bool g_quiet = false;
#define LOGI(...) do { if (!g_quiet) { fprintf(stderr, "[henjou] " __VA_ARGS__); fputc('\n', stderr); } } while (0)

static std::string dir_of(const std::string& p) { const size_t k = p.find_last_of('/'); return k == std::string::npos ? std::string("./") : p.substr(0, k + 1); }
static std::string join(const std::string& dir, const std::string& rel) { return (!rel.empty() && rel[0] == '/') ? rel : dir + rel; }

// Camera of frame `time` (renderer.h:1145-1169): animated branch through the camera node's TRS, static branch un-normalised
static void camera_basis(const RenderOption& o, const HostScene& sc, float time, hj_frame_params& fp) {
    fp.camera_f = (float)(2.0 / std::tan(o.camera_fov));
    float pos[3], dir[3], up[3], right[3];
    if (o.camera_animation_id != -1 && o.allow_camera_animation) {
        const NodeAnimation& a = sc.animations[(size_t)o.camera_animation_id];
        const Affine ap = a.affine(time), ad = a.rotate_only(time);
        const float p4[4] = {o.camera_position[0], o.camera_position[1], o.camera_position[2], 1.0f};
        const float d4[4] = {o.camera_direction[0], o.camera_direction[1], o.camera_direction[2], 0.0f};
        const float u4[4] = {0, 1, 0, 0};
        float tp[4], td[4], tu[4];
        affine_mul4(ap, p4, tp); affine_mul4(ad, d4, td); affine_mul4(ad, u4, tu);
        for (int k = 0; k < 3; k++) { pos[k] = tp[k]; dir[k] = td[k]; up[k] = tu[k]; }
        const float cx = dir[1] * up[2] - dir[2] * up[1], cy = dir[2] * up[0] - dir[0] * up[2], cz = dir[0] * up[1] - dir[1] * up[0];
        const float inv = 1.0f / sqrtf(cx * cx + cy * cy + cz * cz);      // sutil normalize
        right[0] = cx * inv; right[1] = cy * inv; right[2] = cz * inv;
    } else {
        for (int k = 0; k < 3; k++) { pos[k] = o.camera_position[k]; dir[k] = o.camera_direction[k]; }
        const float yup[3] = {0, 1, 0};
        right[0] = dir[1] * yup[2] - dir[2] * yup[1]; right[1] = dir[2] * yup[0] - dir[0] * yup[2]; right[2] = dir[0] * yup[1] - dir[1] * yup[0];
        up[0] = right[1] * dir[2] - right[2] * dir[1]; up[1] = right[2] * dir[0] - right[0] * dir[2]; up[2] = right[0] * dir[1] - right[1] * dir[0];
    }
    memcpy(fp.camera_pos, pos, 12); memcpy(fp.camera_dir, dir, 12); memcpy(fp.camera_up, up, 12); memcpy(fp.camera_right, right, 12);
}

// updateIASMatrix (renderer.h:257-291): instance-to-world rows and the rows of the inverse
static void instance_matrices(const HostScene& sc, float time, std::vector<float>& xf, std::vector<float>& inv) {
    const size_t n = sc.n_instances();
    xf.resize(12 * n); inv.resize(12 * n);
    for (size_t i = 0; i < n; i++) {
        const Affine a = sc.animations[sc.instance_node[i]].affine(time);
        memcpy(&xf[12 * i], a.v, 48);
        affine_inverse_rows(a, &inv[12 * i]);
    }
}

// --dump-scene: what the loaders produced, as named arrays (tests/test_host_loader.py reads it back)
static void dump_section(FILE* f, const char* name, char dtype, const void* data, uint64_t count, size_t elem) {
    char nm[24] = {0};
    strncpy(nm, name, 23);
    fwrite(nm, 1, 24, f); fwrite(&dtype, 1, 1, f); fwrite(&count, 8, 1, f);
    if (count) fwrite(data, elem, count, f);
}
static void dump_scene(const std::string& path, const RenderOption& o, const HostScene& sc, int f0, int f1) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    fwrite("HJSCENE1", 1, 8, f);
    dump_section(f, "vertices", 'f', sc.vertices.data(), sc.vertices.size(), 4);
    dump_section(f, "normals", 'f', sc.normals.data(), sc.normals.size(), 4);
    dump_section(f, "texcoords", 'f', sc.texcoords.data(), sc.texcoords.size(), 4);
    dump_section(f, "indices", 'u', sc.indices.data(), sc.indices.size(), 4);
    dump_section(f, "material_ids", 'u', sc.material_ids.data(), sc.material_ids.size(), 4);
    dump_section(f, "prim_offsets", 'u', sc.prim_offsets.data(), sc.prim_offsets.size(), 4);
    dump_section(f, "light_prim_ids", 'u', sc.light_prim_ids.data(), sc.light_prim_ids.size(), 4);
    dump_section(f, "light_prim_emission", 'f', sc.light_prim_emission.data(), sc.light_prim_emission.size(), 4);
    dump_section(f, "materials", 'b', sc.materials.data(), sc.materials.size() * sizeof(hj_material), 1);
    dump_section(f, "ibl", 'f', sc.ibl.data(), sc.ibl.size(), 4);
    for (size_t i = 0; i < sc.textures.size(); i++) {
        char nm[24]; snprintf(nm, sizeof nm, "tex%zu", i);
        const int32_t hdr[3] = {sc.textures[i].width, sc.textures[i].height, sc.textures[i].srgb};
        char nh[40]; snprintf(nh, sizeof nh, "tex%zu_whs", i);
        dump_section(f, nh, 'i', hdr, 3, 4);
        dump_section(f, nm, 'b', sc.textures[i].rgba.data(), sc.textures[i].rgba.size(), 1);
    }
    const float opt[8] = {(float)o.image_width, (float)o.image_height, (float)o.max_spp, (float)o.fps, o.camera_fov, (float)o.camera_animation_id, o.IBL_intensity, 0.0f};
    dump_section(f, "options", 'f', opt, 8, 4);
    for (int fr = f0; fr < f1; fr++) {          // per frame: matrices, inverse rows, camera
        const float time = fr / float(o.fps);
        std::vector<float> xf, inv;
        instance_matrices(sc, time, xf, inv);
        hj_frame_params fp{};
        camera_basis(o, sc, time, fp);
        char nm[24];
        snprintf(nm, sizeof nm, "xf_%d", fr); dump_section(f, nm, 'f', xf.data(), xf.size(), 4);
        snprintf(nm, sizeof nm, "inv_%d", fr); dump_section(f, nm, 'f', inv.data(), inv.size(), 4);
        const float cam[13] = {fp.camera_pos[0], fp.camera_pos[1], fp.camera_pos[2], fp.camera_dir[0], fp.camera_dir[1], fp.camera_dir[2],
                               fp.camera_up[0], fp.camera_up[1], fp.camera_up[2], fp.camera_right[0], fp.camera_right[1], fp.camera_right[2], fp.camera_f};
        snprintf(nm, sizeof nm, "cam_%d", fr); dump_section(f, nm, 'f', cam, 13, 4);
    }
    fclose(f);
}

#define HJ_CHECK(ctx, call) do { const int rc_ = (call); if (rc_ != HJ_OK) { \
    fprintf(stderr, "henjou_b200_cli: %s failed (%d): %s\n", #call, rc_, hj_last_error(ctx)); return 2; } } while (0)

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv) {
    std::string option_path = "./render_option.json", out_dir = "./", dump_path;
    int device = 0, spp_override = -1, f_begin = -1, f_end = -1, w_override = -1, h_override = -1;
    bool rebuild = false, enforce_limit = false;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto need = [&](int n) { if (i + n >= argc) { fprintf(stderr, "henjou_b200_cli: %s needs %d argument(s)\n", a.c_str(), n); exit(1); } };
        if (a == "--device") { need(1); device = atoi(argv[++i]); }
        else if (a == "--spp") { need(1); spp_override = atoi(argv[++i]); }
        else if (a == "--frames") { need(2); f_begin = atoi(argv[++i]); f_end = atoi(argv[++i]); }
        else if (a == "--width") { need(1); w_override = atoi(argv[++i]); }
        else if (a == "--height") { need(1); h_override = atoi(argv[++i]); }
        else if (a == "--out-dir") { need(1); out_dir = argv[++i]; if (out_dir.back() != '/') out_dir += '/'; }
        else if (a == "--dump-scene") { need(1); dump_path = argv[++i]; }
        else if (a == "--rebuild") rebuild = true;
        else if (a == "--enforce-time-limit") enforce_limit = true;
        else if (a == "--quiet") g_quiet = true;
        else if (a == "--help" || a == "-h") { printf("usage: henjou_b200_cli [render_option.json] [--device N] [--spp N] [--frames BEGIN END] [--width W --height H] [--out-dir DIR] [--dump-scene FILE] [--rebuild] [--enforce-time-limit] [--quiet]\n"); return 0; }
        else if (a[0] != '-') option_path = a;
        else { fprintf(stderr, "henjou_b200_cli: unknown option %s\n", a.c_str()); return 1; }
    }
    RenderOption opt; HostScene sc;
    const std::string base = dir_of(option_path);
    try {
        const double t0 = now_s();
        opt = load_render_option(option_path, base + "fps.txt");
        if (spp_override > 0) opt.max_spp = (unsigned)spp_override;
        if (f_begin >= 0) { opt.start_frame = f_begin; opt.end_frame = f_end; }
        if (w_override > 0) opt.image_width = w_override;
        if (h_override > 0) opt.image_height = h_override;
        LOGI("image %dx%d, %u spp, frames [%d,%d) at %u fps, mode %s", opt.image_width, opt.image_height, opt.max_spp, opt.start_frame, opt.end_frame, opt.fps, opt.render_mode.c_str());
        if (opt.render_mode != "Default")
            LOGI("Render_mode %s: the OptiX AI denoiser is unavailable on this box; hj_post's a-trous filter (albedo + normal guided) stands in", opt.render_mode.c_str());
        load_gltf(join(base, opt.gltf_path), opt.gltf_name, opt, sc);
        // setSky (renderer.h:802-811): the HDR image, or the constant default when use_IBL is off / the file is missing
        sc.ibl = {opt.scene_sky_default[0], opt.scene_sky_default[1], opt.scene_sky_default[2], 0.0f}; sc.ibl_w = sc.ibl_h = 1;
        if (opt.use_IBL) {
            try { sc.ibl = hdr_read(join(base, opt.IBL_path), sc.ibl_w, sc.ibl_h); }
            catch (const std::exception& e) { LOGI("IBL not loaded (%s): constant sky", e.what()); }
        }
        // setLUT (renderer.h:854-898)
        try {
            Image8 im = png_read(join(base, opt.LUT_path));
            sc.lut.width = im.width; sc.lut.height = im.height; sc.lut.rgba = std::move(im.rgba); sc.has_lut = true;
        } catch (const std::exception& e) { LOGI("thin-film LUT not loaded (%s): thin-film F0 is 0", e.what()); }
        LOGI("scene: %u triangles, %u instances, %zu materials, %zu textures, %zu emissive triangles, loaded in %.3f s",
             sc.n_triangles(), sc.n_instances(), sc.materials.size(), sc.textures.size(), sc.light_prim_ids.size(), now_s() - t0);
        if (!dump_path.empty()) { dump_scene(dump_path, opt, sc, opt.start_frame, opt.end_frame); LOGI("scene dumped to %s", dump_path.c_str()); return 0; }
    } catch (const std::exception& e) {
        fprintf(stderr, "henjou_b200_cli: %s\n", e.what());
        return 1;
    }
    if (sc.n_triangles() == 0) { fprintf(stderr, "henjou_b200_cli: the scene has no triangles\n"); return 1; }

    hj_ctx* ctx = nullptr;
    if (hj_create(&ctx, device) != HJ_OK) { fprintf(stderr, "henjou_b200_cli: %s\n", hj_last_error(nullptr)); return 2; }
    std::vector<hj_texture> texs(sc.textures.size());
    for (size_t i = 0; i < texs.size(); i++) texs[i] = hj_texture{sc.textures[i].rgba.data(), sc.textures[i].width, sc.textures[i].height, sc.textures[i].srgb};
    hj_scene_desc d{};
    d.vertices = sc.vertices.data(); d.normals = sc.normals.data(); d.texcoords = sc.texcoords.data();
    d.indices = sc.indices.data(); d.material_ids = sc.material_ids.data(); d.prim_offsets = sc.prim_offsets.data();
    d.n_vertices = (uint32_t)(sc.vertices.size() / 3); d.n_triangles = sc.n_triangles(); d.n_instances = sc.n_instances();
    d.materials = sc.materials.data(); d.n_materials = (uint32_t)sc.materials.size();
    d.textures = texs.data(); d.n_textures = (uint32_t)texs.size();
    d.light_prim_ids = sc.light_prim_ids.data(); d.light_prim_emission = sc.light_prim_emission.data(); d.n_light_prims = (uint32_t)sc.light_prim_ids.size();
    d.ibl_rgba32f = sc.ibl.data(); d.ibl_width = sc.ibl_w; d.ibl_height = sc.ibl_h;
    if (sc.has_lut) { d.lut_rgba8 = sc.lut.rgba.data(); d.lut_width = sc.lut.width; d.lut_height = sc.lut.height; }
    HJ_CHECK(ctx, hj_scene_upload(ctx, &d));

    const double t_start = now_s();
    // renderer.h:1089-1115: UpScale2X renders at half the resolution; Default passes the colour AOV through (blendFactor 1);
    // every other mode (Denoise, Debug) runs the HDR denoiser
    const bool upscale = opt.render_mode == "DenoiseUpScale2X";
    const int post_mode = upscale ? HJ_POST_DENOISE_UPSCALE2X : (opt.render_mode == "Default" ? HJ_POST_PASSTHROUGH : HJ_POST_DENOISE);
    const uint32_t in_w = upscale ? (uint32_t)opt.image_width / 2u : (uint32_t)opt.image_width, in_h = upscale ? (uint32_t)opt.image_height / 2u : (uint32_t)opt.image_height;
    const uint32_t out_w = upscale ? 2u * in_w : in_w, out_h = upscale ? 2u * in_h : in_h;
    if (in_w == 0 || in_h == 0) { fprintf(stderr, "henjou_b200_cli: image too small for Render_mode %s\n", opt.render_mode.c_str()); hj_destroy(ctx); return 1; }
    // the PNG of frame N is deflated and written by a worker while the GPU renders frame N+1 (two image buffers in turn)
    std::vector<uint8_t> rgba_buf[2] = {std::vector<uint8_t>((size_t)out_w * out_h * 4), std::vector<uint8_t>((size_t)out_w * out_h * 4)};
    std::future<std::string> writer;
    auto join_writer = [&]() -> bool {
        if (!writer.valid()) return true;
        const std::string err = writer.get();
        if (!err.empty()) fprintf(stderr, "henjou_b200_cli: %s\n", err.c_str());
        return err.empty();
    };
    int n_done = 0;
    bool first = true;
    // The acceleration structure of frame N+1 is updated on the library's build stream while frame N is shaded
    // (hj_accel_build_async right after hj_render_async): the reference does updateIASMatrix at the top of each iteration
    // (renderer.h:1133) and stalls on it.
    auto enqueue_accel = [&](int frame, bool blocking) -> int {
        const float time = frame / float(opt.fps);                                // :1128
        std::vector<float> xf, inv;
        instance_matrices(sc, time, xf, inv);                                     // updateIASMatrix, :1133
        const int mode = (first || rebuild) ? HJ_ACCEL_BUILD : HJ_ACCEL_REFIT;
        return blocking ? hj_accel_build(ctx, xf.data(), inv.data(), sc.n_instances(), mode)
                        : hj_accel_build_async(ctx, xf.data(), inv.data(), sc.n_instances(), mode);
    };
    HJ_CHECK(ctx, enqueue_accel(opt.start_frame, true));
    first = false;
    for (int frame = opt.start_frame; frame < opt.end_frame; frame++) {          // renderer.h:1126
        const float time = frame / float(opt.fps);                                // :1128
        hj_frame_params fp{};
        fp.width = in_w; fp.height = in_h;
        fp.spp = opt.max_spp; fp.sample_begin = 0; fp.sample_end = opt.max_spp; fp.frame = (uint32_t)frame;      // :1186-1187
        camera_basis(opt, sc, time, fp);
        fp.ibl_intensity = opt.IBL_intensity;
        fp.integrator = opt.integrator; fp.max_depth = opt.max_depth; fp.tonemap = opt.tonemap; fp.pixel_jitter = 1;
        HJ_CHECK(ctx, hj_render_async(ctx, &fp));                                 // optixLaunch, :1241
        if (frame + 1 < opt.end_frame) HJ_CHECK(ctx, enqueue_accel(frame + 1, false));   // next frame's update under this frame's shading
        HJ_CHECK(ctx, hj_sync(ctx));                                              // CUDA_SYNC_CHECK, :1242
        HJ_CHECK(ctx, hj_post(ctx, post_mode));                                   // denosier_manager.denoise(), :1262-1271
        std::vector<uint8_t>& rgba = rgba_buf[n_done & 1];                        // its previous user (frame N-2) was joined below
        HJ_CHECK(ctx, hj_readback(ctx, HJ_AOV_OUTPUT_RGBA8, rgba.data(), rgba.size()));   // float4ConvertColor(AOV_Output), :1281
        char num[16]; snprintf(num, sizeof num, "%03d", frame);                   // :1291-1301
        const std::string name = out_dir + opt.image_name + "_" + num + ".png";
        if (!join_writer()) { hj_destroy(ctx); return 1; }
        writer = std::async(std::launch::async, [name, &rgba, out_w, out_h]() -> std::string {
            try { png_write(name, rgba.data(), (int)out_w, (int)out_h, true); } catch (const std::exception& e) { return e.what(); }
            return std::string();
        });
        n_done++;
        hj_stats st{}; hj_get_stats(ctx, &st);
        const double el = now_s() - t_start;
        LOGI("frame %d: accel %.2f ms, render %.2f ms, post %.2f ms, %.1f Mrays/s -> %s   (time limit: %.1f s / %.1f s)", frame, st.ms_accel, st.ms_render, st.ms_post,
             (st.rays_extend + st.rays_shadow) / (st.ms_render * 1e3), name.c_str(), el, opt.time_limit * 60.0f);
        // the check the reference carries commented out (renderer.h:1247-1254): stop once the sequence is over its budget
        if (enforce_limit && el > opt.time_limit * 60.0f) { LOGI("Over Time Limit: %d of %d frames written", n_done, opt.end_frame - opt.start_frame); break; }
    }
    const bool ok = join_writer();
    hj_destroy(ctx);
    return ok ? 0 : 1;
}
