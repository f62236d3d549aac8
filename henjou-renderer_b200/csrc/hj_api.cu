// hj_api.cu — the C ABI of libhenjou_b200.so (include/henjou_b200.h): context, scene upload, option handling,
// read-back. Replaces the host<->device plumbing of the reference's Renderer
// (/root/reference/include/renderer/renderer.h: cpySceneDataToDevice :197-255, textureBind :740-800,
// setSky :802-851, setLUT :854-898, updateIASMatrix :257-291, launch :1229-1242, AOV read-back :1279-1283).
// No exception crosses the boundary; there is no CPU path: without an sm_100 device hj_create fails.
#include "hj_ctx.h"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <new>

static std::string g_create_err;

#define HJ_REQUIRE(ctx, cond, code, msg)            \
    do { if (!(cond)) { (ctx)->err = (msg); return (code); } } while (0)

template <class T>
static int upload(hj_ctx* c, DevBuf& b, const T* src, size_t n) {
    HJ_CUDA(c, b.reserve(std::max<size_t>(n, 1) * sizeof(T)));
    if (n) HJ_CUDA(c, cudaMemcpyAsync(b.p, src, n * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    return HJ_OK;
}

// tri_inst[t] = the instance whose [prim_offsets[i], prim_offsets[i+1]) contains triangle t (prim_offsets strictly increasing)
__global__ void k_tri_inst(const uint32_t* __restrict__ prim_offsets, uint32_t n_inst, uint32_t n_tris, uint32_t* __restrict__ tri_inst) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tris) return;
    uint32_t lo = 0, hi = n_inst;                 // last i with prim_offsets[i] <= t
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (prim_offsets[mid] <= t) lo = mid; else hi = mid; }
    tri_inst[t] = lo;
}

// Sky sampling table (frame parameter sky_sampling): row y, sequentially: running sum of luminance(texel) x sin(pi (y + 1/2) / H); then
// the running sum of the row totals. One thread per row / one thread: the SAME float addition order as the oracle's loops.
__global__ void k_sky_rows(const float4* __restrict__ ibl, int W, int H, float* __restrict__ row_cdf) {
    const int y = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= H) return;
    const float st = hj_sinf(M_PIf * ((float)y + 0.5f) / (float)H);
    float run = 0.0f;
    for (int x = 0; x < W; x++) {
        const float4 c = ibl[(size_t)y * W + x];
        const float lum = 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z;
        run += fmaxf(lum, 0.0f) * st;
        row_cdf[(size_t)y * W + x] = run;
    }
}
__global__ void k_sky_marginal(const float* __restrict__ row_cdf, int W, int H, float* __restrict__ marg) {
    float run = 0.0f;
    for (int y = 0; y < H; y++) { run += row_cdf[(size_t)y * W + W - 1]; marg[y] = run; }
}

__global__ void k_libm_eval(int fn, const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, uint64_t n) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = a[i], y = b ? b[i] : 0.0f;
    float r = 0.0f;
    switch (fn) {
        case 0: r = hj_sinf(x); break;
        case 1: r = hj_cosf(x); break;
        case 2: r = hj_acosf(x); break;
        case 3: r = hj_logf(x); break;
        case 4: r = hj_expf(x); break;
        case 5: r = hj_powf(x, y); break;
        default: r = hj_atan2f(x, y); break;
    }
    out[i] = r;
}

extern "C" {

int hj_abi_version(void) { return HJ_ABI_VERSION; }

const char* hj_last_error(const hj_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int hj_create(hj_ctx** out, int device) {
    if (!out) { g_create_err = "hj_create: out is NULL"; return HJ_ERR_INVALID; }
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) { g_create_err = std::string("hj_create: no CUDA device (") + cudaGetErrorString(e) + "); there is no CPU fallback"; return HJ_ERR_CUDA; }
    if (device < 0 || device >= n) { g_create_err = "hj_create: device index out of range"; return HJ_ERR_INVALID; }
    cudaDeviceProp p;
    if ((e = cudaGetDeviceProperties(&p, device)) != cudaSuccess) { g_create_err = cudaGetErrorString(e); return HJ_ERR_CUDA; }
    if (p.major != 10) {
        char b[200]; snprintf(b, sizeof b, "hj_create: device %d is sm_%d%d; this library contains sm_100a code only", device, p.major, p.minor);
        g_create_err = b; return HJ_ERR_CUDA;
    }
    if ((e = cudaSetDevice(device)) != cudaSuccess) { g_create_err = cudaGetErrorString(e); return HJ_ERR_CUDA; }
    hj_ctx* c = new (std::nothrow) hj_ctx();
    if (!c) { g_create_err = "hj_create: out of host memory"; return HJ_ERR_OOM; }
    c->device = device;
    if (const char* v = getenv("HJ_VERIFY_ACCEL")) c->opt_verify_accel = atoi(v) != 0;    // the test suite sets it
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess || (e = cudaEventCreate(&c->ev0)) != cudaSuccess ||
        (e = cudaEventCreate(&c->ev1)) != cudaSuccess || (e = cudaEventCreateWithFlags(&c->ev_scene, cudaEventDisableTiming)) != cudaSuccess) {
        g_create_err = cudaGetErrorString(e); delete c; return HJ_ERR_CUDA;
    }
    *out = c;
    return HJ_OK;
}

void hj_comm_destroy_internal(hj_ctx* c);

void hj_destroy(hj_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->bvh.build_stream) cudaStreamSynchronize(c->bvh.build_stream);     // a build enqueued by hj_accel_build_async may still be writing the buffers freed below
    for (WavefrontState& w : c->wf) if (w.stream) cudaStreamSynchronize(w.stream);
    hj_comm_destroy_internal(c);
    DevBuf* bufs[] = {&c->d_vertices, &c->d_normals, &c->d_texcoords, &c->d_indices, &c->d_material_ids, &c->d_prim_offsets, &c->d_tri_inst,
                      &c->d_materials, &c->d_textures, &c->d_texels, &c->d_srgb, &c->d_light_ids, &c->d_light_em,
                      &c->d_ibl, &c->d_lut, &c->d_sky_rows, &c->d_sky_marg,
                      &c->bvh.keys, &c->bvh.keys_alt, &c->bvh.vals, &c->bvh.vals_alt, &c->bvh.block_hist, &c->bvh.digit_total, &c->bvh.status, &c->bvh.blo, &c->bvh.bhi, &c->bvh.left,
                      &c->bvh.right, &c->bvh.parent, &c->bvh.cnt, &c->bvh.flags, &c->bvh.p_cid[0], &c->bvh.p_cid[1], &c->bvh.p_lo[0], &c->bvh.p_lo[1],
                      &c->bvh.p_hi[0], &c->bvh.p_hi[1], &c->bvh.p_nn, &c->bvh.p_keep, &c->bvh.p_bcount, &c->bvh.bounds, &c->bvh.queue_a, &c->bvh.queue_b,
                      &c->bvh.counters, &c->bvh.verr,
                      &c->aov_color, &c->aov_albedo, &c->aov_normal, &c->aov_rgba8, &c->aov_accum, &c->aov_path_rad, &c->aov_path_ids,
                      &c->aov_output, &c->aov_output_rgba8, &c->post_tmp[0], &c->post_tmp[1]};
    for (DevBuf* b : bufs) b->release();
    for (DevBuf& b : c->trace_scratch) b.release();
    hj_bvh_release(c);
    for (WavefrontState& w : c->wf) {
        for (DevBuf* b : w.all) if (b) b->release();
        if (w.stream) cudaStreamDestroy(w.stream);
        if (w.ev_acc) cudaEventDestroy(w.ev_acc);
    }
    if (c->ev_frame) cudaEventDestroy(c->ev_frame);
    if (c->ev_post0) cudaEventDestroy(c->ev_post0);
    if (c->ev_post1) cudaEventDestroy(c->ev_post1);
    for (cudaEvent_t e : c->ev_pool) cudaEventDestroy(e);
    cudaEventDestroy(c->ev0); cudaEventDestroy(c->ev1); cudaEventDestroy(c->ev_scene);
    cudaStreamDestroy(c->stream);
    delete c;
}

int hj_set_option(hj_ctx* c, const char* key, int64_t value) {
    if (!c || !key) return HJ_ERR_INVALID;
    if (!strcmp(key, "count_fetches")) c->opt_count_fetches = value != 0;
    else if (!strcmp(key, "time_stages")) c->opt_time_stages = value != 0;
    else if (!strcmp(key, "paths_per_pass")) c->opt_paths_per_pass = value;
    else if (!strcmp(key, "trace_blocks_per_sm")) { HJ_REQUIRE(c, value >= 0 && value <= 32, HJ_ERR_INVALID, "hj_set_option: trace_blocks_per_sm must be 0 (auto) .. 32"); c->opt_trace_blocks_per_sm = (int)value; }
    else if (!strcmp(key, "pipelines")) { HJ_REQUIRE(c, value >= 0 && value <= HJ_MAX_PIPES, HJ_ERR_INVALID, "hj_set_option: pipelines must be 0 (auto) .. 8"); c->opt_pipelines = (int)value; }
    else if (!strcmp(key, "builder")) { HJ_REQUIRE(c, value == 0 || value == 1, HJ_ERR_INVALID, "hj_set_option: builder must be 0 (LBVH) or 1 (PLOC)"); c->opt_builder = (int)value; c->bvh.have_topology = false; }
    else if (!strcmp(key, "leaf_max")) { HJ_REQUIRE(c, value >= 1 && value <= 3, HJ_ERR_INVALID, "hj_set_option: leaf_max must be 1..3"); c->opt_leaf_max = (int)value; }
    else if (!strcmp(key, "leaf_split")) c->opt_leaf_split = value != 0;
    else if (!strcmp(key, "bands")) { HJ_REQUIRE(c, value >= 0 && value <= 64, HJ_ERR_INVALID, "hj_set_option: bands must be 0 (auto) .. 64"); c->opt_bands = (int)value; }
    else if (!strcmp(key, "ploc_tail")) { c->opt_ploc_tail = value != 0; c->bvh.have_topology = false; }
    else if (!strcmp(key, "verify_accel")) c->opt_verify_accel = value != 0;
    else if (!strcmp(key, "nan_repair")) c->opt_nan_repair = value != 0;
    else if (!strcmp(key, "morton_bits")) { HJ_REQUIRE(c, value >= 4 && value <= 21, HJ_ERR_INVALID, "hj_set_option: morton_bits (per axis) must be 4..21"); c->opt_morton_bits = (int)value; c->bvh.have_topology = false; }
    else if (!strcmp(key, "mis_emitter_first")) c->opt_mis_emitter_first = value != 0;
    else if (!strcmp(key, "denoise_iterations")) { HJ_REQUIRE(c, value >= 0 && value <= 8, HJ_ERR_INVALID, "hj_set_option: denoise_iterations must be 0..8"); c->opt_denoise_iterations = (int)value; }
    else if (!strncmp(key, "denoise_sigma_", 14)) {
        HJ_REQUIRE(c, value >= 1 && value <= 1000000, HJ_ERR_INVALID, "hj_set_option: denoise sigmas are in 1/1000, 1..1000000");
        if (!strcmp(key + 14, "color_milli")) c->opt_denoise_sigma_color_milli = (int)value;
        else if (!strcmp(key + 14, "normal_milli")) c->opt_denoise_sigma_normal_milli = (int)value;
        else if (!strcmp(key + 14, "albedo_milli")) c->opt_denoise_sigma_albedo_milli = (int)value;
        else { c->err = std::string("hj_set_option: unknown option '") + key + "'"; return HJ_ERR_INVALID; }
    }
    else { c->err = std::string("hj_set_option: unknown option '") + key + "'"; return HJ_ERR_INVALID; }
    return HJ_OK;
}

static float srgb_to_linear(float v) { return (v <= 0.04045f) ? v / 12.92f : hjm_powf((v + 0.055f) / 1.055f, 2.4f); }   // hj_libm.h: same table as the oracle's

int hj_scene_upload_impl(hj_ctx* c, const hj_scene_desc* d, bool geometry_from_host);

int hj_scene_upload(hj_ctx* c, const hj_scene_desc* d) { return hj_scene_upload_impl(c, d, true); }

// geometry_from_host = false (hj_comm_scene_upload on a rank that is not the root): the five big SceneData arrays are only
// allocated; their contents arrive over NVLink. Everything small (offsets, materials, textures, lights, sky, LUT) is uploaded as usual.
int hj_scene_upload_impl(hj_ctx* c, const hj_scene_desc* d, bool geometry_from_host) {
    if (!c) return HJ_ERR_INVALID;
    HJ_REQUIRE(c, d && d->prim_offsets && (!geometry_from_host || (d->vertices && d->indices && d->material_ids)), HJ_ERR_INVALID, "hj_scene_upload: missing geometry arrays");
    HJ_REQUIRE(c, d->n_instances > 0 && d->n_triangles > 0 && d->n_vertices > 0, HJ_ERR_INVALID, "hj_scene_upload: empty scene");
    HJ_REQUIRE(c, d->n_materials > 0 && d->materials, HJ_ERR_INVALID, "hj_scene_upload: no materials");
    HJ_REQUIRE(c, d->prim_offsets[0] == 0, HJ_ERR_INVALID, "hj_scene_upload: prim_offsets[0] must be 0");
    for (uint32_t i = 1; i < d->n_instances; i++)
        HJ_REQUIRE(c, d->prim_offsets[i] > d->prim_offsets[i - 1] && d->prim_offsets[i] < d->n_triangles, HJ_ERR_INVALID, "hj_scene_upload: prim_offsets must be strictly increasing and < n_triangles");
    for (uint32_t i = 0; i < d->n_light_prims; i++)
        HJ_REQUIRE(c, d->light_prim_ids && d->light_prim_emission && d->light_prim_ids[i] < d->n_triangles, HJ_ERR_INVALID, "hj_scene_upload: bad light primitive list");
    HJ_CUDA(c, cudaSetDevice(c->device));
    (void)hj_bvh_finish(c);                                  // a build in flight reads the arrays replaced below
    HJ_CUDA(c, cudaStreamSynchronize(c->stream));
    c->have_scene = c->have_accel = c->have_frame = c->have_output = false;
    c->bvh.have_topology = false;
    int rc;
    const size_t nv = d->n_vertices, nt = d->n_triangles;
    if (geometry_from_host) {
        if ((rc = upload(c, c->d_vertices, d->vertices, 3 * nv))) return rc;
        {   // normals default to +y and uvs to 0 when absent (the loader always provides them, gltfloader.h:1460-1482)
            std::vector<float> tmp;
            if (!d->normals) { tmp.assign(3 * nv, 0.0f); for (size_t i = 0; i < nv; i++) tmp[3 * i + 1] = 1.0f; }
            if ((rc = upload(c, c->d_normals, d->normals ? d->normals : tmp.data(), 3 * nv))) return rc;
            if (!d->normals) HJ_CUDA(c, cudaStreamSynchronize(c->stream));          // tmp is reused / freed below
            if (!d->texcoords) tmp.assign(2 * nv, 0.0f);
            if ((rc = upload(c, c->d_texcoords, d->texcoords ? d->texcoords : tmp.data(), 2 * nv))) return rc;
            if (!d->texcoords) HJ_CUDA(c, cudaStreamSynchronize(c->stream));
        }
        if ((rc = upload(c, c->d_indices, d->indices, 3 * nt))) return rc;
        if ((rc = upload(c, c->d_material_ids, d->material_ids, nt))) return rc;
    } else {
        HJ_CUDA(c, c->d_vertices.reserve(3 * nv * 4)); HJ_CUDA(c, c->d_normals.reserve(3 * nv * 4)); HJ_CUDA(c, c->d_texcoords.reserve(2 * nv * 4));
        HJ_CUDA(c, c->d_indices.reserve(3 * nt * 4)); HJ_CUDA(c, c->d_material_ids.reserve(nt * 4));
    }
    if ((rc = upload(c, c->d_prim_offsets, d->prim_offsets, (size_t)d->n_instances))) return rc;
    // owning instance per triangle: derived on the device from prim_offsets (no host pass, no extra upload)
    HJ_CUDA(c, c->d_tri_inst.reserve(nt * sizeof(uint32_t)));
    k_tri_inst<<<(uint32_t)((nt + 255) / 256), 256, 0, c->stream>>>(c->d_prim_offsets.as<uint32_t>(), d->n_instances, d->n_triangles, c->d_tri_inst.as<uint32_t>());
    // range checks of the two big index arrays run on the host while the copies above are in flight (they come out of
    // page-locked memory when the caller used hj_host_pin); a violation leaves the context without a scene
    if (geometry_from_host) {
        uint32_t max_index = 0, max_material = 0;
        const uint32_t* idx = d->indices; const size_t ni = 3 * nt;
        for (size_t i = 0; i < ni; i++) max_index = idx[i] > max_index ? idx[i] : max_index;
        for (size_t i = 0; i < nt; i++) max_material = d->material_ids[i] > max_material ? d->material_ids[i] : max_material;
        if (max_index >= d->n_vertices || max_material >= d->n_materials) {
            cudaStreamSynchronize(c->stream);
            c->err = max_index >= d->n_vertices ? "hj_scene_upload: vertex index out of range" : "hj_scene_upload: material id out of range";
            return HJ_ERR_INVALID;
        }
    }
    std::vector<DevMaterial> mats(d->n_materials);
    for (uint32_t i = 0; i < d->n_materials; i++) {
        const hj_material& m = d->materials[i]; DevMaterial& o = mats[i];
        for (int k = 0; k < 3; k++) { o.basecolor[k] = m.basecolor[k]; o.emission[k] = m.emission[k]; }
        o.basecolor_tex = m.basecolor_tex; o.is_light = m.is_light; o.metallic = m.metallic; o.roughness = m.roughness;
        o.roughness_tex = m.roughness_tex; o.sheen = m.sheen; o.clearcoat = m.clearcoat; o.ior = m.ior; o.transmission = m.transmission;
        o.flags = (m.ideal_specular ? 1 : 0) | (m.is_thinfilm ? 2 : 0);
    }
    if ((rc = upload(c, c->d_materials, mats.data(), mats.size()))) return rc;
    // textures: one texel pool + descriptors
    size_t texels = 0;
    for (uint32_t i = 0; i < d->n_textures; i++) {
        HJ_REQUIRE(c, d->textures && d->textures[i].rgba8 && d->textures[i].width > 0 && d->textures[i].height > 0, HJ_ERR_INVALID, "hj_scene_upload: bad texture");
        texels += (size_t)d->textures[i].width * d->textures[i].height;
    }
    HJ_CUDA(c, c->d_texels.reserve(std::max<size_t>(texels, 1) * 4));
    std::vector<DevTexture> texd(std::max<uint32_t>(d->n_textures, 1));
    size_t off = 0;
    for (uint32_t i = 0; i < d->n_textures; i++) {
        const size_t n = (size_t)d->textures[i].width * d->textures[i].height;
        HJ_CUDA(c, cudaMemcpyAsync(c->d_texels.as<uchar4>() + off, d->textures[i].rgba8, n * 4, cudaMemcpyHostToDevice, c->stream));
        texd[i].px = c->d_texels.as<uchar4>() + off; texd[i].w = d->textures[i].width; texd[i].h = d->textures[i].height; texd[i].srgb = d->textures[i].srgb;
        off += n;
    }
    if ((rc = upload(c, c->d_textures, texd.data(), texd.size()))) return rc;
    float table[256];
    for (int i = 0; i < 256; i++) table[i] = srgb_to_linear((float)i / 255.0f);   // sRGB EOTF per texel (renderer.h:785)
    if ((rc = upload(c, c->d_srgb, table, (size_t)256))) return rc;
    if ((rc = upload(c, c->d_light_ids, d->light_prim_ids, (size_t)d->n_light_prims))) return rc;
    if ((rc = upload(c, c->d_light_em, d->light_prim_emission, 3 * (size_t)d->n_light_prims))) return rc;
    const float black[4] = {0, 0, 0, 0};
    int iw = 1, ih = 1;
    if (d->ibl_rgba32f && d->ibl_width > 0 && d->ibl_height > 0) { iw = d->ibl_width; ih = d->ibl_height; if ((rc = upload(c, c->d_ibl, d->ibl_rgba32f, 4 * (size_t)iw * ih))) return rc; }
    else if ((rc = upload(c, c->d_ibl, black, (size_t)4))) return rc;
    HJ_CUDA(c, c->d_sky_rows.reserve((size_t)iw * ih * 4)); HJ_CUDA(c, c->d_sky_marg.reserve((size_t)ih * 4));
    k_sky_rows<<<(ih + 63) / 64, 64, 0, c->stream>>>(c->d_ibl.as<float4>(), iw, ih, c->d_sky_rows.as<float>());
    k_sky_marginal<<<1, 1, 0, c->stream>>>(c->d_sky_rows.as<float>(), iw, ih, c->d_sky_marg.as<float>());
    const bool has_lut = d->lut_rgba8 && d->lut_width > 0 && d->lut_height > 0;
    if (has_lut) { if ((rc = upload(c, c->d_lut, d->lut_rgba8, 4 * (size_t)d->lut_width * d->lut_height))) return rc; }
    HJ_CUDA(c, cudaEventRecord(c->ev_scene, c->stream));
    HJ_CUDA(c, cudaStreamSynchronize(c->stream));
    DevScene& S = c->S;
    S = DevScene{};
    S.vertices = c->d_vertices.as<float>(); S.normals = c->d_normals.as<float>(); S.texcoords = c->d_texcoords.as<float>();
    S.indices = c->d_indices.as<uint32_t>(); S.material_ids = c->d_material_ids.as<uint32_t>(); S.prim_offsets = c->d_prim_offsets.as<uint32_t>();
    S.tri_inst = c->d_tri_inst.as<uint32_t>();
    S.n_vertices = d->n_vertices; S.n_tris = d->n_triangles; S.n_instances = d->n_instances; S.n_materials = d->n_materials;
    S.n_textures = d->n_textures; S.n_lights = d->n_light_prims;
    S.materials = c->d_materials.as<DevMaterial>(); S.textures = c->d_textures.as<DevTexture>(); S.srgb_table = c->d_srgb.as<float>();
    S.light_prim_ids = c->d_light_ids.as<uint32_t>(); S.light_prim_emission = c->d_light_em.as<float>();
    S.ibl = c->d_ibl.as<float4>(); S.ibl_w = iw; S.ibl_h = ih;
    S.sky_row_cdf = c->d_sky_rows.as<float>(); S.sky_marg = c->d_sky_marg.as<float>();
    S.lut = c->d_lut.as<uchar4>(); S.lut_w = has_lut ? d->lut_width : 0; S.lut_h = has_lut ? d->lut_height : 0; S.has_lut = has_lut;
    c->have_scene = true;
    return HJ_OK;
}

static int accel_args_ok(hj_ctx* c, uint32_t n_instances, int mode, const char* who) {
    HJ_REQUIRE(c, c->have_scene, HJ_ERR_NO_SCENE, std::string(who) + ": no scene uploaded");
    HJ_REQUIRE(c, n_instances == c->S.n_instances, HJ_ERR_INVALID, std::string(who) + ": n_instances differs from the uploaded scene");
    HJ_REQUIRE(c, mode == HJ_ACCEL_BUILD || mode == HJ_ACCEL_REFIT, HJ_ERR_INVALID, std::string(who) + ": unknown mode");
    return HJ_OK;
}

int hj_accel_build_async(hj_ctx* c, const float* xf, const float* inv_xf, uint32_t n_instances, int mode) {
    if (!c) return HJ_ERR_INVALID;
    int rc = accel_args_ok(c, n_instances, mode, "hj_accel_build_async");
    if (rc) return rc;
    HJ_CUDA(c, cudaSetDevice(c->device));
    if ((rc = hj_bvh_finish(c))) return rc;               // at most one build in flight
    return hj_bvh_enqueue(c, xf, inv_xf, mode);
}

int hj_accel_wait(hj_ctx* c) {
    if (!c) return HJ_ERR_INVALID;
    HJ_CUDA(c, cudaSetDevice(c->device));
    return hj_bvh_finish(c);
}

int hj_accel_build(hj_ctx* c, const float* xf, const float* inv_xf, uint32_t n_instances, int mode) {
    int rc = hj_accel_build_async(c, xf, inv_xf, n_instances, mode);
    return rc ? rc : hj_bvh_finish(c);
}

int hj_render_async(hj_ctx* c, const hj_frame_params* fp) {
    if (!c || !fp) return HJ_ERR_INVALID;
    HJ_REQUIRE(c, c->have_scene, HJ_ERR_NO_SCENE, "hj_render: no scene uploaded");
    HJ_CUDA(c, cudaSetDevice(c->device));
    { int rc = hj_bvh_finish(c); if (rc) return rc; }     // a build enqueued by hj_accel_build_async: validate it and flip to its buffers
    HJ_REQUIRE(c, c->have_accel, HJ_ERR_NO_ACCEL, "hj_render: hj_accel_build has not been called");
    return hj_wavefront_render(c, fp);
}

int hj_sync(hj_ctx* c) {
    if (!c) return HJ_ERR_INVALID;
    HJ_CUDA(c, cudaSetDevice(c->device));
    HJ_CUDA(c, cudaStreamSynchronize(c->stream));
    HJ_CUDA(c, cudaGetLastError());
    return c->have_frame ? hj_wavefront_stats(c) : HJ_OK;
}

int hj_render(hj_ctx* c, const hj_frame_params* fp) {
    int rc = hj_render_async(c, fp);
    return rc ? rc : hj_sync(c);
}

static int aov_info(hj_ctx* c, int aov, void** p, size_t* bytes) {
    const size_t wh = (size_t)c->fp.width * c->fp.height, ns = c->fp.sample_end - c->fp.sample_begin;
    switch (aov) {
        case HJ_AOV_COLOR: *p = c->aov_color.p; *bytes = wh * 16; return HJ_OK;
        case HJ_AOV_ALBEDO: *p = c->aov_albedo.p; *bytes = wh * 16; return HJ_OK;
        case HJ_AOV_NORMAL: *p = c->aov_normal.p; *bytes = wh * 16; return HJ_OK;
        case HJ_AOV_RGBA8: *p = c->aov_rgba8.p; *bytes = wh * 4; return HJ_OK;
        case HJ_AOV_ACCUM: *p = c->aov_accum.p; *bytes = wh * 16; return HJ_OK;
        case HJ_AOV_PATH_RADIANCE: *p = c->fp.debug_paths ? c->aov_path_rad.p : nullptr; *bytes = ns * wh * 16; return HJ_OK;
        case HJ_AOV_PATH_IDS: *p = c->fp.debug_paths ? c->aov_path_ids.p : nullptr; *bytes = ns * wh * 8; return HJ_OK;
        case HJ_AOV_OUTPUT: *p = c->have_output ? c->aov_output.p : nullptr; *bytes = (size_t)c->out_w * c->out_h * 16; return HJ_OK;
        case HJ_AOV_OUTPUT_RGBA8: *p = c->have_output ? c->aov_output_rgba8.p : nullptr; *bytes = (size_t)c->out_w * c->out_h * 4; return HJ_OK;
    }
    return HJ_ERR_INVALID;
}

int hj_readback(hj_ctx* c, int aov, void* dst, size_t bytes) {
    if (!c || !dst) return HJ_ERR_INVALID;
    HJ_REQUIRE(c, c->have_frame, HJ_ERR_NO_FRAME, "hj_readback: nothing rendered yet");
    void* p = nullptr; size_t want = 0;
    HJ_REQUIRE(c, aov_info(c, aov, &p, &want) == HJ_OK, HJ_ERR_INVALID, "hj_readback: unknown AOV");
    HJ_REQUIRE(c, p != nullptr, HJ_ERR_INVALID, "hj_readback: AOV not recorded (debug_paths was 0 / hj_post not called for this frame)");
    HJ_REQUIRE(c, bytes == want, HJ_ERR_INVALID, "hj_readback: size mismatch");
    HJ_CUDA(c, cudaSetDevice(c->device));
    HJ_CUDA(c, cudaMemcpyAsync(dst, p, bytes, cudaMemcpyDeviceToHost, c->stream));
    HJ_CUDA(c, cudaStreamSynchronize(c->stream));
    return HJ_OK;
}

int hj_post(hj_ctx* c, int mode) {
    if (!c) return HJ_ERR_INVALID;
    HJ_REQUIRE(c, c->have_frame, HJ_ERR_NO_FRAME, "hj_post: nothing rendered yet");
    HJ_REQUIRE(c, mode >= HJ_POST_PASSTHROUGH && mode <= HJ_POST_DENOISE_UPSCALE2X, HJ_ERR_INVALID, "hj_post: unknown mode");
    HJ_CUDA(c, cudaSetDevice(c->device));
    if (!c->ev_post0) { HJ_CUDA(c, cudaEventCreate(&c->ev_post0)); HJ_CUDA(c, cudaEventCreate(&c->ev_post1)); }
    return hj_post_run(c, mode);
}

int hj_host_pin(hj_ctx* c, void* ptr, size_t bytes) {
    if (!c || !ptr || !bytes) return HJ_ERR_INVALID;
    HJ_CUDA(c, cudaSetDevice(c->device));
    HJ_CUDA(c, cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
    return HJ_OK;
}

int hj_host_unpin(hj_ctx* c, void* ptr) {
    if (!c || !ptr) return HJ_ERR_INVALID;
    HJ_CUDA(c, cudaSetDevice(c->device));
    HJ_CUDA(c, cudaHostUnregister(ptr));
    return HJ_OK;
}

void* hj_device_ptr(hj_ctx* c, int aov) {
    if (!c || !c->have_frame) return nullptr;
    void* p = nullptr; size_t b = 0;
    return aov_info(c, aov, &p, &b) == HJ_OK ? p : nullptr;
}

int hj_finalize(hj_ctx* c) {
    if (!c) return HJ_ERR_INVALID;
    HJ_REQUIRE(c, c->have_frame, HJ_ERR_NO_FRAME, "hj_finalize: nothing rendered yet");
    HJ_CUDA(c, cudaSetDevice(c->device));
    int rc = hj_wavefront_finalize(c);
    if (rc) return rc;
    HJ_CUDA(c, cudaStreamSynchronize(c->stream));
    return HJ_OK;
}

int hj_libm_eval(hj_ctx* c, int fn, const float* a, const float* b, float* out, uint64_t n) {
    if (!c || !a || !out) return HJ_ERR_INVALID;
    HJ_REQUIRE(c, fn >= 0 && fn <= 6, HJ_ERR_INVALID, "hj_libm_eval: fn must be 0..6");
    HJ_REQUIRE(c, fn < 5 || b, HJ_ERR_INVALID, "hj_libm_eval: binary function needs b");
    if (n == 0) return HJ_OK;
    HJ_CUDA(c, cudaSetDevice(c->device));
    DevBuf da, db, dout;
    HJ_CUDA(c, da.reserve(n * 4)); HJ_CUDA(c, dout.reserve(n * 4));
    HJ_CUDA(c, cudaMemcpyAsync(da.p, a, n * 4, cudaMemcpyHostToDevice, c->stream));
    if (b) { HJ_CUDA(c, db.reserve(n * 4)); HJ_CUDA(c, cudaMemcpyAsync(db.p, b, n * 4, cudaMemcpyHostToDevice, c->stream)); }
    k_libm_eval<<<(uint32_t)((n + 255) / 256), 256, 0, c->stream>>>(fn, da.as<float>(), b ? db.as<float>() : nullptr, dout.as<float>(), n);
    HJ_CUDA(c, cudaMemcpyAsync(out, dout.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
    HJ_CUDA(c, cudaStreamSynchronize(c->stream));
    HJ_CUDA(c, cudaGetLastError());
    da.release(); db.release(); dout.release();
    return HJ_OK;
}

int hj_get_stats(hj_ctx* c, hj_stats* out) {
    if (!c || !out) return HJ_ERR_INVALID;
    *out = c->stats;
    return HJ_OK;
}

int hj_trace(hj_ctx* c, const float* rays, uint64_t n, int any_hit, uint32_t* prim, float* tuv) {
    if (!c || (n && !rays)) return HJ_ERR_INVALID;
    HJ_CUDA(c, cudaSetDevice(c->device));
    { int rc = hj_bvh_finish(c); if (rc) return rc; }
    HJ_REQUIRE(c, c->have_accel, HJ_ERR_NO_ACCEL, "hj_trace: hj_accel_build has not been called");
    return hj_trace_batch(c, rays, n, any_hit, prim, tuv);
}

}  // extern "C"
