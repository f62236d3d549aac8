// hj_comm.cu — multi-GPU: sample-slice data parallelism with a replicated scene. The reference is
// single-GPU (one stream, no collectives: /root/reference/include/renderer/renderer.h:1077-1078); the one
// exchange step this build adds is a sum-reduction of the float4 accumulation buffer per frame, over NCCL on
// NVLink 5 / NVSwitch. NCCL is bound at run time (dlopen) so that a process which already loaded a NCCL
// (e.g. through torch) shares it instead of mixing two copies.
#include "hj_ctx.h"
#include <dlfcn.h>
#include <nccl.h>

struct NcclApi {
    void* lib = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclBroadcast) Broadcast = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
};

static NcclApi* load_nccl(std::string& err) {
    static NcclApi api; static bool tried = false;
    if (tried) return api.CommInitRank ? &api : nullptr;
    tried = true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);        // a copy this process already has
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { err = std::string("NCCL not found: ") + dlerror(); return nullptr; }
    api.lib = h;
    api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))dlsym(h, "ncclCommInitRank");
    api.AllReduce = (decltype(api.AllReduce))dlsym(h, "ncclAllReduce");
    api.Broadcast = (decltype(api.Broadcast))dlsym(h, "ncclBroadcast");
    api.CommDestroy = (decltype(api.CommDestroy))dlsym(h, "ncclCommDestroy");
    api.GetErrorString = (decltype(api.GetErrorString))dlsym(h, "ncclGetErrorString");
    if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.CommDestroy) { err = "NCCL symbols missing"; api.CommInitRank = nullptr; return nullptr; }
    return &api;
}

static std::string g_comm_err;

extern "C" {

int hj_comm_unique_id(void* unique_id_128) {
    if (!unique_id_128) return HJ_ERR_INVALID;
    NcclApi* n = load_nccl(g_comm_err);
    if (!n) return HJ_ERR_COMM;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if (n->GetUniqueId(&id) != ncclSuccess) return HJ_ERR_COMM;
    memcpy(unique_id_128, &id, 128);
    return HJ_OK;
}

int hj_comm_init(hj_ctx* c, const void* unique_id_128, int rank, int nranks) {
    if (!c || !unique_id_128 || nranks < 1 || rank < 0 || rank >= nranks) return HJ_ERR_INVALID;
    NcclApi* n = load_nccl(c->err);
    if (!n) return HJ_ERR_COMM;
    HJ_CUDA(c, cudaSetDevice(c->device));
    ncclUniqueId id; memcpy(&id, unique_id_128, 128);
    ncclComm_t comm = nullptr;
    ncclResult_t r = n->CommInitRank(&comm, nranks, id, rank);
    if (r != ncclSuccess) { c->err = std::string("ncclCommInitRank: ") + (n->GetErrorString ? n->GetErrorString(r) : "error"); return HJ_ERR_COMM; }
    c->nccl = n; c->comm = comm; c->rank = rank; c->nranks = nranks;
    return HJ_OK;
}

// all-reduce (sum) of the float4 accumulation buffer of the last frame, then mean/tonemap/quantise on every rank
int hj_comm_reduce(hj_ctx* c) {
    if (!c) return HJ_ERR_INVALID;
    if (!c->have_frame) { c->err = "hj_comm_reduce: nothing rendered yet"; return HJ_ERR_NO_FRAME; }
    HJ_CUDA(c, cudaSetDevice(c->device));
    if (c->comm && c->nranks > 1 && !c->frame_reduced) {    // in place: summing twice would scale the frame by nranks
        const size_t count = 4 * (size_t)c->fp.width * c->fp.height;
        ncclResult_t r = c->nccl->AllReduce(c->aov_accum.p, c->aov_accum.p, count, ncclFloat32, ncclSum, (ncclComm_t)c->comm, c->stream);
        if (r != ncclSuccess) { c->err = std::string("ncclAllReduce: ") + (c->nccl->GetErrorString ? c->nccl->GetErrorString(r) : "error"); return HJ_ERR_COMM; }
        c->frame_reduced = true;
    }
    int rc = hj_wavefront_finalize(c);
    if (rc) return rc;
    HJ_CUDA(c, cudaStreamSynchronize(c->stream));
    return HJ_OK;
}

// Scene upload for a replicated scene: ONE host-to-device copy of the geometry (on `root`), then the five big SceneData arrays go
// to the other ranks with ncclBroadcast over NVLink / NVSwitch instead of one PCIe copy per rank through the same host
// (r01: at 8 ranks the eight concurrent 207 MB - 1.1 GB uploads were what kept the end-to-end figure at 0.85 of linear).
int hj_scene_upload_impl(hj_ctx* c, const hj_scene_desc* d, bool geometry_from_host);
int hj_comm_scene_upload(hj_ctx* c, const hj_scene_desc* d, int root) {
    if (!c || !d) return HJ_ERR_INVALID;
    if (!c->comm || c->nranks <= 1) return hj_scene_upload_impl(c, d, true);
    if (root < 0 || root >= c->nranks) { c->err = "hj_comm_scene_upload: root out of range"; return HJ_ERR_INVALID; }
    if (!c->nccl->Broadcast) { c->err = "hj_comm_scene_upload: ncclBroadcast not found"; return HJ_ERR_COMM; }
    HJ_CUDA(c, cudaSetDevice(c->device));
    // every rank must reach the broadcasts, so a rank whose own part failed still takes part and reports afterwards
    const int rc = hj_scene_upload_impl(c, d, c->rank == root);
    const std::string first_err = c->err;
    if (rc != HJ_OK && (c->d_vertices.cap < 12 * (size_t)d->n_vertices || c->d_indices.cap < 12 * (size_t)d->n_triangles)) return rc;   // nothing to receive into
    const size_t nv = d->n_vertices, nt = d->n_triangles;
    struct { DevBuf* b; size_t bytes; } parts[5] = {{&c->d_vertices, 12 * nv}, {&c->d_normals, 12 * nv}, {&c->d_texcoords, 8 * nv},
                                                    {&c->d_indices, 12 * nt}, {&c->d_material_ids, 4 * nt}};
    for (auto& p : parts) {
        ncclResult_t r = c->nccl->Broadcast(p.b->p, p.b->p, p.bytes, ncclChar, root, (ncclComm_t)c->comm, c->stream);
        if (r != ncclSuccess) { c->err = std::string("ncclBroadcast: ") + (c->nccl->GetErrorString ? c->nccl->GetErrorString(r) : "error"); return HJ_ERR_COMM; }
    }
    HJ_CUDA(c, cudaEventRecord(c->ev_scene, c->stream));
    HJ_CUDA(c, cudaStreamSynchronize(c->stream));
    if (rc != HJ_OK) { c->err = first_err; c->have_scene = false; }
    return rc;
}

void hj_comm_destroy_internal(hj_ctx* c) {
    if (c->comm && c->nccl) c->nccl->CommDestroy((ncclComm_t)c->comm);
    c->comm = nullptr;
}

}  // extern "C"
