// C ABI of libb200clip (declared in include/b200clip.h): handle, weight upload, and the orchestration of
// the kernels in preprocess.cu / gemm_tc.cu / vit_ops.cu / topk.cu / classify.cu.
#include "../../include/b200clip.h"
#include "common.cuh"
#include "kernels.h"

#include <dlfcn.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <set>
#include <string>
#include <type_traits>
#include <vector>

namespace b200 {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
int fail(int code, const std::string& msg) { g_last_error = msg; return code; }

// fp32 [rows, cols] -> bf16 rows of pitch ld (destination padding, if any, stays zero)
__global__ void f32_to_bf16_rows(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long rows, int cols, int ld) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * cols) return;
    const long long r = i / cols; const int c = (int)(i - r * cols);
    dst[r * ld + c] = __float2bfloat16(src[i]);
}

struct Layer {
    __nv_bfloat16 *w_qkv = nullptr, *w_o = nullptr, *w_fc1 = nullptr, *w_fc2 = nullptr;
    float *b_qkv = nullptr, *b_o = nullptr, *b_fc1 = nullptr, *b_fc2 = nullptr;
    float *ln1_g = nullptr, *ln1_b = nullptr, *ln2_g = nullptr, *ln2_b = nullptr;
};

// N4: the CLIP text tower (label-embedding bootstrap, general_clip/clip_model.py:145-172) living next to the image tower on the
// same handle.  Same Layer weights and the same GEMM / LayerNorm kernels; attention is causal; pooling takes the EOT token.
struct TextTower {
    b200clip_text_config cfg{};
    std::vector<void*> allocs;
    float *tok = nullptr, *pos = nullptr, *final_g = nullptr, *final_b = nullptr;
    __nv_bfloat16* w_proj = nullptr;
    std::vector<Layer> layers;
    std::set<std::string> loaded;
    // activations for max_batch sequences
    float *xres = nullptr, *eot = nullptr, *emb_raw = nullptr;
    __nv_bfloat16 *hbuf = nullptr, *qkv = nullptr, *mlp = nullptr, *eot_ln = nullptr;
    size_t want_tensors() const { return 5 + 16 * (size_t)cfg.layers; }
};

}  // namespace b200

using namespace b200;

struct b200clip_handle {
    b200clip_config cfg{};
    std::mutex mu;
    int n = 0, S = 0, kpad = 0;
    std::vector<void*> allocs;
    // weights
    __nv_bfloat16 *w_patch = nullptr, *w_proj = nullptr;
    float *cls = nullptr, *pos = nullptr, *pre_g = nullptr, *pre_b = nullptr, *post_g = nullptr, *post_b = nullptr;
    std::vector<Layer> layers;
    std::set<std::string> loaded;
    float* upload_tmp = nullptr; size_t upload_cap = 0;
    // activations workspace
    __nv_bfloat16 *patches = nullptr, *hbuf = nullptr, *qkv = nullptr, *mlp = nullptr, *cls_ln = nullptr;
    float *patch_out = nullptr, *xres = nullptr, *emb_raw = nullptr;
    // last layer on the CLS rows only (B200_CLS_LAST=0 computes every row, as the earlier layers do)
    __nv_bfloat16 *q_cls = nullptr, *attn_cls = nullptr, *mlp_cls = nullptr;
    bool cls_last = true;
    // host-buffer path staging
    uint8_t* img_dev = nullptr; size_t img_cap = 0;
    float* emb_dev = nullptr; size_t emb_cap = 0;
    // classify label sets
    float* labels[2] = {nullptr, nullptr}; int labels_n[2] = {0, 0}, labels_d[2] = {0, 0};
    // search scratch
    void* topk_scratch = nullptr; size_t topk_cap = 0;
    void* shard_buf = nullptr; size_t shard_cap = 0;      // b200clip_search_sharded: own keys | gathered keys | own base | gathered bases
    PreprocessState* pre = nullptr;
    JpegState* jpeg = nullptr;          // created on first decode
    TextTower* text = nullptr;          // b200clip_text_init
    DeviceStatus* status = nullptr;
    int64_t launches = 0;
    // The workspace (activations, descriptors, decode arenas, search scratch) is one per handle.  The mutex orders the host side of
    // two calls; when they name different streams, the later one also has to wait on the device for the earlier one's kernels.
    cudaStream_t ws_stream = nullptr; cudaEvent_t ws_done = nullptr; bool ws_used = false;
    int ws_begin(cudaStream_t st) {
        if (ws_used && st != ws_stream) B200_CUDA_OK(cudaStreamWaitEvent(st, ws_done, 0));
        return B200_OK;
    }
    int ws_end(cudaStream_t st) {
        if (!ws_done) B200_CUDA_OK(cudaEventCreateWithFlags(&ws_done, cudaEventDisableTiming));
        B200_CUDA_OK(cudaEventRecord(ws_done, st));
        ws_stream = st; ws_used = true;
        return B200_OK;
    }
    // optional per-kernel-class timing (bench.py roofline): CUDA events recorded around each launch
    bool prof = false;
    std::vector<cudaEvent_t> ev_pool; size_t ev_used = 0;
    struct ProfRec { int cls; cudaEvent_t a, b; };
    std::vector<ProfRec> prof_recs;
    cudaEvent_t prof_event() {
        if (ev_used == ev_pool.size()) { cudaEvent_t e; cudaEventCreate(&e); ev_pool.push_back(e); }
        return ev_pool[ev_used++];
    }

    // The image tower's weights live in a block that b200clip_clone()d handles share (freed with the last of them); everything else
    // (activations, descriptors, scratch) is per handle.
    struct WeightBlock {
        std::vector<void*> allocs;
        ~WeightBlock() { for (void* q : allocs) cudaFree(q); }
    };
    std::shared_ptr<WeightBlock> wblock;
    bool alloc_into_wblock = false;
    bool is_clone = false;

    template <class T> int alloc(T** p, size_t count, bool zero = false) {
        void* q = nullptr;
        B200_CUDA_OK(cudaMalloc(&q, count * sizeof(T) + 256));
        if (zero) B200_CUDA_OK(cudaMemset(q, 0, count * sizeof(T) + 256));
        if (alloc_into_wblock) wblock->allocs.push_back(q); else allocs.push_back(q);
        *p = reinterpret_cast<T*>(q);
        return B200_OK;
    }
};

namespace b200 {
// Programmatic dependent launch pays when a pass is a chain of short kernels (batch 1: 0.87 -> 0.69 ms for ViT-B/32); at batch 512
// it measured neutral to slightly negative (ViT-L/14: +1.7 % time), so the encoder turns it on per call for small passes only.
// B200_PDL=0 / 1 forces it off / on.
static thread_local bool tl_pdl_small_pass = false;
bool pdl_enabled() {
    static const int mode = [] { const char* e = getenv("B200_PDL"); return e ? (atoi(e) != 0 ? 1 : 0) : -1; }();
    return mode >= 0 ? mode == 1 : tl_pdl_small_pass;
}
struct PdlScope {
    bool prev;
    explicit PdlScope(bool on) : prev(tl_pdl_small_pass) { tl_pdl_small_pass = on; }
    ~PdlScope() { tl_pdl_small_pass = prev; }
};
}  // namespace b200

static int set_device(b200clip_handle* h) {
    B200_REQUIRE(h != nullptr, "null handle");
    B200_CUDA_OK(cudaSetDevice(h->cfg.device));
    return B200_OK;
}

extern "C" const char* b200clip_last_error(void) { return g_last_error.c_str(); }

static int alloc_activations(b200clip_handle* h);
static int require_encoder(b200clip_handle* h);

extern "C" int b200clip_create(const b200clip_config* cfg, b200clip_handle** out) {
    B200_REQUIRE(cfg && out, "create: null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(B200CLIP_EDEVICE, "no CUDA device: libb200clip has no CPU fallback");
    B200_REQUIRE(cfg->device >= 0 && cfg->device < ndev, "create: device ordinal out of range");
    cudaDeviceProp prop;
    B200_CUDA_OK(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10) return fail(B200CLIP_EDEVICE, std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) +
                                                        ", libb200clip is built for sm_100a only");
    B200_REQUIRE(cfg->image_size > 0 && cfg->image_size % 2 == 0, "create: image_size must be positive and even");
    for (int c = 0; c < 3; ++c) B200_REQUIRE(cfg->std[c] != 0.f, "create: std must be non-zero");
    B200_CUDA_OK(cudaSetDevice(cfg->device));
    auto* h = new b200clip_handle();
    h->cfg = *cfg;
    int rc = B200_OK;
    h->pre = preprocess_state_create(cfg->mean, cfg->std, &rc);
    if (!h->pre) { delete h; return rc; }
    auto bail = [&](int code) { b200clip_destroy(h); return code; };
    if ((rc = h->alloc(&h->status, 1, true))) return bail(rc);
    if (cfg->layers > 0) {
        const int d = cfg->width, P = cfg->patch, L = cfg->layers, mlp = cfg->mlp_dim, D = cfg->embed_dim;
        if (!(P > 0 && P % 2 == 0 && cfg->image_size % P == 0)) return bail(fail(B200CLIP_EINVAL, "create: image_size must be a multiple of an even patch size"));
        if (!(d % 128 == 0 && d <= 1024 && d == cfg->heads * 64)) return bail(fail(B200CLIP_EINVAL, "create: width must be heads*64, a multiple of 128, at most 1024"));
        if (!(mlp % 64 == 0 && D > 0 && cfg->max_batch > 0)) return bail(fail(B200CLIP_EINVAL, "create: mlp_dim must be a multiple of 64; embed_dim and max_batch positive"));
        h->n = (cfg->image_size / P) * (cfg->image_size / P);
        h->S = h->n + 1;
        h->kpad = (3 * P * P + 63) / 64 * 64;
        const size_t Bm = cfg->max_batch, M = Bm * h->S, Mp = Bm * h->n;
        h->layers.resize(L);
#define A_(call) if ((rc = (call))) return bail(rc)
        h->wblock = std::make_shared<b200clip_handle::WeightBlock>();
        h->alloc_into_wblock = true;
        A_(h->alloc(&h->w_patch, (size_t)d * h->kpad, true));
        A_(h->alloc(&h->w_proj, (size_t)D * d));
        A_(h->alloc(&h->cls, d)); A_(h->alloc(&h->pos, (size_t)h->S * d));
        A_(h->alloc(&h->pre_g, d)); A_(h->alloc(&h->pre_b, d)); A_(h->alloc(&h->post_g, d)); A_(h->alloc(&h->post_b, d));
        for (auto& l : h->layers) {
            A_(h->alloc(&l.w_qkv, (size_t)3 * d * d)); A_(h->alloc(&l.b_qkv, 3 * d));
            A_(h->alloc(&l.w_o, (size_t)d * d)); A_(h->alloc(&l.b_o, d));
            A_(h->alloc(&l.w_fc1, (size_t)mlp * d)); A_(h->alloc(&l.b_fc1, mlp));
            A_(h->alloc(&l.w_fc2, (size_t)d * mlp)); A_(h->alloc(&l.b_fc2, d));
            A_(h->alloc(&l.ln1_g, d)); A_(h->alloc(&l.ln1_b, d)); A_(h->alloc(&l.ln2_g, d)); A_(h->alloc(&l.ln2_b, d));
        }
        h->alloc_into_wblock = false;
        if ((rc = alloc_activations(h))) return bail(rc);
#undef A_
    }
    *out = h;
    return B200_OK;
}

// activations and per-handle scratch of the image tower (b200clip_create and b200clip_clone)
static int alloc_activations(b200clip_handle* h) {
    const b200clip_config* cfg = &h->cfg;
    const int d = cfg->width, mlp = cfg->mlp_dim, D = cfg->embed_dim;
    const size_t Bm = cfg->max_batch, M = Bm * h->S, Mp = Bm * h->n;
    int rc;
    {
#define A_(call) if ((rc = (call))) return rc
        A_(h->alloc(&h->patches, Mp * h->kpad, true));     // K padding must read as zero
        A_(h->alloc(&h->patch_out, Mp * d));
        A_(h->alloc(&h->xres, M * d));
        A_(h->alloc(&h->hbuf, M * d));
        A_(h->alloc(&h->qkv, M * 3 * d));
        A_(h->alloc(&h->mlp, M * mlp));
        A_(h->alloc(&h->cls_ln, Bm * d));
        A_(h->alloc(&h->emb_raw, Bm * D));
        { const char* e = getenv("B200_CLS_LAST"); h->cls_last = e ? atoi(e) != 0 : true; }
        A_(h->alloc(&h->q_cls, Bm * d)); A_(h->alloc(&h->attn_cls, Bm * d)); A_(h->alloc(&h->mlp_cls, Bm * mlp));
#undef A_
    }
    return B200_OK;
}

extern "C" int b200clip_clone(b200clip_handle* src, b200clip_handle** out) {
    int rc = set_device(src); if (rc) return rc;
    B200_REQUIRE(out != nullptr, "clone: null argument");
    std::lock_guard<std::mutex> lk(src->mu);
    B200_REQUIRE(src->cfg.layers > 0, "clone: the handle has no image tower");
    if ((rc = require_encoder(src))) return rc;
    auto* h = new b200clip_handle();
    h->cfg = src->cfg;
    h->is_clone = true;
    h->pre = preprocess_state_create(h->cfg.mean, h->cfg.std, &rc);
    if (!h->pre) { delete h; return rc; }
    auto bail = [&](int code) { b200clip_destroy(h); return code; };
    if ((rc = h->alloc(&h->status, 1, true))) return bail(rc);
    h->n = src->n; h->S = src->S; h->kpad = src->kpad;
    // the weights: same device memory, kept alive by the shared block
    h->wblock = src->wblock;
    h->w_patch = src->w_patch; h->w_proj = src->w_proj; h->cls = src->cls; h->pos = src->pos;
    h->pre_g = src->pre_g; h->pre_b = src->pre_b; h->post_g = src->post_g; h->post_b = src->post_b;
    h->layers = src->layers;
    h->loaded = src->loaded;
    if ((rc = alloc_activations(h))) return bail(rc);
    *out = h;
    return B200_OK;
}

extern "C" void b200clip_destroy(b200clip_handle* h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    cudaDeviceSynchronize();
    for (void* p : h->allocs) cudaFree(p);
    for (cudaEvent_t e : h->ev_pool) cudaEventDestroy(e);
    if (h->ws_done) cudaEventDestroy(h->ws_done);
    cudaFree(h->upload_tmp); cudaFree(h->img_dev); cudaFree(h->emb_dev); cudaFree(h->topk_scratch); cudaFree(h->shard_buf);
    cudaFree(h->labels[0]); cudaFree(h->labels[1]);
    if (h->pre) preprocess_state_destroy(h->pre);
    jpeg_state_destroy(h->jpeg);
    if (h->text) { for (void* p : h->text->allocs) cudaFree(p); delete h->text; }
    delete h;
}

extern "C" int b200clip_sync(b200clip_handle* h, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream));
    DeviceStatus st;
    B200_CUDA_OK(cudaMemcpy(&st, h->status, sizeof(st), cudaMemcpyDeviceToHost));
    if (st.code != 0) {
        cudaMemset(h->status, 0, sizeof(DeviceStatus));
        return fail(B200CLIP_EKERNEL, "device watchdog: pipeline barrier never completed (site " + std::to_string(st.where) + ")");
    }
    return B200_OK;
}

extern "C" int64_t b200clip_launch_count(b200clip_handle* h) { return h ? h->launches : 0; }

extern "C" int b200clip_profile(b200clip_handle* h, int32_t enable) {
    B200_REQUIRE(h != nullptr, "null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    h->prof = enable != 0;          // records accumulate until b200clip_profile_read
    return B200_OK;
}

extern "C" int b200clip_profile_read(b200clip_handle* h, float* ms_by_class, int64_t* launches_by_class, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(ms_by_class && launches_by_class, "profile_read: null buffer");
    std::lock_guard<std::mutex> lk(h->mu);
    B200_CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream));
    for (int c = 0; c < 4; ++c) { ms_by_class[c] = 0.f; launches_by_class[c] = 0; }
    for (const auto& r : h->prof_recs) {
        float ms = 0.f;
        B200_CUDA_OK(cudaEventElapsedTime(&ms, r.a, r.b));
        ms_by_class[r.cls] += ms; launches_by_class[r.cls] += 1;
    }
    h->prof_recs.clear(); h->ev_used = 0;
    return B200_OK;
}

// ------------------------------------------------------------------------------------------ weights
static int upload_f32(b200clip_handle* h, const float* host, size_t count, float** dev_tmp) {
    if (count > h->upload_cap) {
        cudaFree(h->upload_tmp); h->upload_tmp = nullptr; h->upload_cap = 0;
        B200_CUDA_OK(cudaMalloc(&h->upload_tmp, count * sizeof(float)));
        h->upload_cap = count;
    }
    B200_CUDA_OK(cudaMemcpy(h->upload_tmp, host, count * sizeof(float), cudaMemcpyHostToDevice));
    *dev_tmp = h->upload_tmp;
    return B200_OK;
}
static int put_bf16(b200clip_handle* h, const float* host, long long rows, int cols, __nv_bfloat16* dst, int ld) {
    float* tmp; int rc = upload_f32(h, host, (size_t)rows * cols, &tmp); if (rc) return rc;
    const long long total = rows * cols;
    f32_to_bf16_rows<<<(unsigned)((total + 255) / 256), 256>>>(tmp, dst, rows, cols, ld);
    B200_CUDA_OK(cudaGetLastError());
    B200_CUDA_OK(cudaDeviceSynchronize());
    return B200_OK;
}
static int put_f32(const float* host, size_t count, float* dst) {
    B200_CUDA_OK(cudaMemcpy(dst, host, count * sizeof(float), cudaMemcpyHostToDevice));
    return B200_OK;
}

// One tensor of a transformer layer (the image and the text tower share the layout): `t` is the key below "...layers.N."
static int load_layer_tensor(b200clip_handle* h, Layer& l, const std::string& t, const std::string& nm, const float* data, size_t count,
                             const int64_t* shape, int ndim, int d, int mlp) {
    int rc = B200_OK;
    auto want = [&](std::initializer_list<int64_t> s) {
        if ((int)s.size() != ndim) return false;
        int i = 0; for (int64_t v : s) if (shape[i++] != v) return false;
        return true;
    };
#define SHAPE_(...) B200_REQUIRE(want({__VA_ARGS__}), "load_weight: wrong shape for " + nm)
        const char* qkvn[3] = {"q_proj", "k_proj", "v_proj"};
        bool done = false;
        for (int j = 0; j < 3 && !done; ++j) {
            if (t == std::string("self_attn.") + qkvn[j] + ".weight") { SHAPE_(d, d); rc = put_bf16(h, data, d, d, l.w_qkv + (size_t)j * d * d, d); done = true; }
            else if (t == std::string("self_attn.") + qkvn[j] + ".bias") { SHAPE_(d); rc = put_f32(data, count, l.b_qkv + j * d); done = true; }
        }
        if (done) {}
        else if (t == "self_attn.out_proj.weight") { SHAPE_(d, d); rc = put_bf16(h, data, d, d, l.w_o, d); }
        else if (t == "self_attn.out_proj.bias") { SHAPE_(d); rc = put_f32(data, count, l.b_o); }
        else if (t == "layer_norm1.weight") { SHAPE_(d); rc = put_f32(data, count, l.ln1_g); }
        else if (t == "layer_norm1.bias") { SHAPE_(d); rc = put_f32(data, count, l.ln1_b); }
        else if (t == "layer_norm2.weight") { SHAPE_(d); rc = put_f32(data, count, l.ln2_g); }
        else if (t == "layer_norm2.bias") { SHAPE_(d); rc = put_f32(data, count, l.ln2_b); }
        else if (t == "mlp.fc1.weight") { SHAPE_(mlp, d); rc = put_bf16(h, data, mlp, d, l.w_fc1, d); }
        else if (t == "mlp.fc1.bias") { SHAPE_(mlp); rc = put_f32(data, count, l.b_fc1); }
        else if (t == "mlp.fc2.weight") { SHAPE_(d, mlp); rc = put_bf16(h, data, d, mlp, l.w_fc2, mlp); }
        else if (t == "mlp.fc2.bias") { SHAPE_(d); rc = put_f32(data, count, l.b_fc2); }
        else return fail(B200CLIP_EINVAL, "load_weight: unknown tensor " + nm);
#undef SHAPE_
    return rc;
}

// text_model.* / text_projection.weight (transformers CLIPTextModelWithProjection keys)
static int load_text_weight(b200clip_handle* h, const std::string& nm, const float* data, const int64_t* shape, int ndim) {
    TextTower* t = h->text;
    B200_REQUIRE(t != nullptr, "load_weight: text tensor but b200clip_text_init was not called");
    const int d = t->cfg.width, mlp = t->cfg.mlp_dim, D = t->cfg.embed_dim, V = t->cfg.vocab_size, ctx = t->cfg.context_length;
    size_t count = 1;
    for (int i = 0; i < ndim; ++i) count *= (size_t)shape[i];
    auto want = [&](std::initializer_list<int64_t> s) {
        if ((int)s.size() != ndim) return false;
        int i = 0; for (int64_t v : s) if (shape[i++] != v) return false;
        return true;
    };
#define SHAPE_(...) B200_REQUIRE(want({__VA_ARGS__}), "load_weight: wrong shape for " + nm)
    int rc = B200_OK;
    const std::string LP = "text_model.encoder.layers.";
    if (nm == "text_model.embeddings.token_embedding.weight") { SHAPE_(V, d); rc = put_f32(data, count, t->tok); }
    else if (nm == "text_model.embeddings.position_embedding.weight") { SHAPE_(ctx, d); rc = put_f32(data, count, t->pos); }
    else if (nm == "text_model.final_layer_norm.weight") { SHAPE_(d); rc = put_f32(data, count, t->final_g); }
    else if (nm == "text_model.final_layer_norm.bias") { SHAPE_(d); rc = put_f32(data, count, t->final_b); }
    else if (nm == "text_projection.weight") { SHAPE_(D, d); rc = put_bf16(h, data, D, d, t->w_proj, d); }
    else if (nm.compare(0, LP.size(), LP) == 0) {
        const size_t dot = nm.find('.', LP.size());
        B200_REQUIRE(dot != std::string::npos, "load_weight: malformed layer key " + nm);
        const int li = std::atoi(nm.substr(LP.size(), dot - LP.size()).c_str());
        B200_REQUIRE(li >= 0 && li < t->cfg.layers, "load_weight: layer index out of range in " + nm);
        rc = load_layer_tensor(h, t->layers[li], nm.substr(dot + 1), nm, data, count, shape, ndim, d, mlp);
    } else return fail(B200CLIP_EINVAL, "load_weight: unknown tensor " + nm);
#undef SHAPE_
    if (rc == B200_OK) t->loaded.insert(nm);
    return rc;
}

extern "C" int b200clip_load_weight(b200clip_handle* h, const char* name, const float* data, const int64_t* shape, int32_t ndim) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(name && data && shape && ndim >= 1 && ndim <= 4, "load_weight: null argument or bad rank");
    B200_REQUIRE(!h->is_clone, "load_weight: a cloned handle shares its source's weights; load them through the source");
    std::lock_guard<std::mutex> lk(h->mu);
    const std::string nm(name);
    if (nm.compare(0, 5, "text_") == 0) return load_text_weight(h, nm, data, shape, ndim);
    B200_REQUIRE(h->cfg.layers > 0, "load_weight: handle was created without an encoder (layers == 0)");
    const int d = h->cfg.width, P = h->cfg.patch, mlp = h->cfg.mlp_dim, D = h->cfg.embed_dim;
    size_t count = 1;
    for (int i = 0; i < ndim; ++i) count *= (size_t)shape[i];
    auto want = [&](std::initializer_list<int64_t> s) {
        if ((int)s.size() != ndim) return false;
        int i = 0; for (int64_t v : s) if (shape[i++] != v) return false;
        return true;
    };
#define SHAPE_(...) B200_REQUIRE(want({__VA_ARGS__}), "load_weight: wrong shape for " + nm)
    const std::string E = "vision_model.embeddings.", LP = "vision_model.encoder.layers.";
    if (nm == E + "class_embedding") { SHAPE_(d); rc = put_f32(data, count, h->cls); }
    else if (nm == E + "patch_embedding.weight") { SHAPE_(d, 3, P, P); rc = put_bf16(h, data, d, 3 * P * P, h->w_patch, h->kpad); }
    else if (nm == E + "position_embedding.weight") { SHAPE_(h->S, d); rc = put_f32(data, count, h->pos); }
    else if (nm == "vision_model.pre_layrnorm.weight") { SHAPE_(d); rc = put_f32(data, count, h->pre_g); }
    else if (nm == "vision_model.pre_layrnorm.bias") { SHAPE_(d); rc = put_f32(data, count, h->pre_b); }
    else if (nm == "vision_model.post_layernorm.weight") { SHAPE_(d); rc = put_f32(data, count, h->post_g); }
    else if (nm == "vision_model.post_layernorm.bias") { SHAPE_(d); rc = put_f32(data, count, h->post_b); }
    else if (nm == "visual_projection.weight") { SHAPE_(D, d); rc = put_bf16(h, data, D, d, h->w_proj, d); }
    else if (nm.compare(0, LP.size(), LP) == 0) {
        const size_t dot = nm.find('.', LP.size());
        B200_REQUIRE(dot != std::string::npos, "load_weight: malformed layer key " + nm);
        const int li = std::atoi(nm.substr(LP.size(), dot - LP.size()).c_str());
        B200_REQUIRE(li >= 0 && li < h->cfg.layers, "load_weight: layer index out of range in " + nm);
        rc = load_layer_tensor(h, h->layers[li], nm.substr(dot + 1), nm, data, count, shape, ndim, d, mlp);
    } else return fail(B200CLIP_EINVAL, "load_weight: unknown tensor " + nm);
#undef SHAPE_
    if (rc == B200_OK) h->loaded.insert(nm);
    return rc;
}

extern "C" int b200clip_finalize_weights(b200clip_handle* h) {
    B200_REQUIRE(h != nullptr, "null handle");
    std::lock_guard<std::mutex> lk(h->mu);
    const size_t want = 8 + 16 * (size_t)h->cfg.layers;
    if (h->text && h->text->loaded.size() != h->text->want_tensors())
        return fail(B200CLIP_ESTATE, "finalize_weights: text tower has " + std::to_string(h->text->loaded.size()) + " of " +
                                     std::to_string(h->text->want_tensors()) + " tensors loaded");
    if (h->cfg.layers == 0 && h->text) { cudaFree(h->upload_tmp); h->upload_tmp = nullptr; h->upload_cap = 0; return B200_OK; }
    if (h->cfg.layers == 0 || h->loaded.size() != want)
        return fail(B200CLIP_ESTATE, "finalize_weights: " + std::to_string(h->loaded.size()) + " of " + std::to_string(want) + " tensors loaded");
    cudaFree(h->upload_tmp); h->upload_tmp = nullptr; h->upload_cap = 0;
    return B200_OK;
}

// ------------------------------------------------------------------------------------------ preprocess / encode
extern "C" int b200clip_preprocess(b200clip_handle* h, const uint8_t* src_dev, const int64_t* offsets, const int32_t* hw, int32_t B,
                                   int32_t mode, int32_t layout, int32_t out_bf16, void* out_dev, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE((reinterpret_cast<uintptr_t>(src_dev) & 15) == 0, "preprocess: src_dev must be 16-byte aligned");
    std::lock_guard<std::mutex> lk(h->mu);
    const int P = h->cfg.patch;
    const int kpad = P > 0 ? (3 * P * P + 63) / 64 * 64 : 0;
    if ((rc = h->ws_begin((cudaStream_t)stream))) return rc;
    rc = preprocess_run(h->pre, src_dev, reinterpret_cast<const long long*>(offsets), hw, B, h->cfg.image_size, mode, layout,
                        out_bf16, P, kpad, out_dev, (cudaStream_t)stream);
    if (rc == B200_OK && B > 0) h->launches += 1;
    if (rc == B200_OK) rc = h->ws_end((cudaStream_t)stream);
    return rc;
}

static int encode_chunk(b200clip_handle* h, const __nv_bfloat16* patches, int B, float* emb_out, cudaStream_t st) {
    const b200clip_config& c = h->cfg;
    const int d = c.width, S = h->S, M = B * S, Mp = B * h->n, mlp = c.mlp_dim, D = c.embed_dim;
    int rc;
    PdlScope pdl(M <= 8192);
#define R_(call) if ((rc = (call))) return rc
    auto timed = [&](int cls, auto&& fn) -> int {
        if (!h->prof) return fn();
        cudaEvent_t a = h->prof_event(), b = h->prof_event();
        cudaEventRecord(a, st);
        const int r = fn();
        cudaEventRecord(b, st);
        h->prof_recs.push_back({cls, a, b});
        return r;
    };
    R_(timed(1, [&] { return gemm_bf16_tc(patches, h->kpad, h->w_patch, h->kpad, Mp, d, h->kpad, nullptr, h->patch_out, d, EPI_F32, 0, h->status, st); }));
    R_(timed(2, [&] { return launch_embed_preln(h->patch_out, h->cls, h->pos, h->pre_g, h->pre_b, h->xres, B, S, d, c.ln_eps, st); }));
    for (size_t li = 0; li < h->layers.size(); ++li) {
        const Layer& l = h->layers[li];
        if (h->cls_last && li + 1 == h->layers.size()) {
            // The tower's output is a function of the CLS row of the last hidden state alone (post-LN and projection are applied to
            // last_hidden_state[:, 0], modeling_clip.py:684-694), and inside a layer rows only meet in the attention.  So the last layer
            // computes keys and values for every token, but the query, the attention output, the out-projection, LN2 and the MLP for
            // the CLS row of each image only: B rows instead of B*S.  The CLS rows go through exactly the arithmetic they always
            // did; what is dropped are rows nothing reads.  (5/6 of the last layer's GEMM work: 6.9 % of a 12-layer tower.)
            const long long cls_pitch = (long long)S * d;                  // CLS rows inside the [B*S, d] buffers
            static const int snake_l = [] { const char* e = getenv("B200_SNAKE"); return e ? atoi(e) : 3; }();
            R_(timed(2, [&] { return launch_layernorm(h->xres, 1, l.ln1_g, l.ln1_b, h->hbuf, M, d, c.ln_eps, st, ((snake_l >> 1) & 1) && li > 0); }));
            R_(timed(1, [&] { return gemm_bf16_tc(h->hbuf, d, l.w_qkv + (size_t)d * d, d, M, 2 * d, d, l.b_qkv + d, h->qkv + d, 3 * d, EPI_BIAS_BF16, 0, h->status, st); }));
            R_(timed(1, [&] { return gemm_bf16_tc(h->hbuf, (int)cls_pitch, l.w_qkv, d, B, d, d, l.b_qkv, h->q_cls, d, EPI_BIAS_BF16, 0, h->status, st); }));
            R_(timed(3, [&] { return launch_cls_attention(h->q_cls, h->qkv, h->attn_cls, B, S, d, c.heads, st); }));
            R_(timed(1, [&] { return gemm_bf16_tc(h->attn_cls, d, l.w_o, d, B, d, d, l.b_o, h->xres, (int)cls_pitch, EPI_BIAS_RES_F32, 0, h->status, st); }));
            R_(timed(2, [&] { return launch_layernorm(h->xres, S, l.ln2_g, l.ln2_b, h->cls_ln, B, d, c.ln_eps, st); }));
            R_(timed(1, [&] { return gemm_bf16_tc(h->cls_ln, d, l.w_fc1, d, B, mlp, d, l.b_fc1, h->mlp_cls, mlp, EPI_BIAS_ACT_BF16, c.activation, h->status, st); }));
            R_(timed(1, [&] { return gemm_bf16_tc(h->mlp_cls, mlp, l.w_fc2, mlp, B, d, mlp, l.b_fc2, h->xres, (int)cls_pitch, EPI_BIAS_RES_F32, 0, h->status, st); }));
            h->launches += 1;                                               // eight kernels instead of the usual seven
            break;
        }
        static const int snake = [] { const char* e = getenv("B200_SNAKE"); return e ? atoi(e) : 3; }();   // bit 0: attention, bit 1: fc2 + LN1
        const int ra = snake & 1, rf = (snake >> 1) & 1;
        R_(timed(2, [&] { return launch_layernorm(h->xres, 1, l.ln1_g, l.ln1_b, h->hbuf, M, d, c.ln_eps, st, rf && li > 0); }));
        R_(timed(1, [&] { return gemm_bf16_tc(h->hbuf, d, l.w_qkv, d, M, 3 * d, d, l.b_qkv, h->qkv, 3 * d, EPI_BIAS_BF16, 0, h->status, st); }));
        R_(timed(3, [&] { return launch_attention_auto(h->qkv, h->hbuf, B, S, d, c.heads, 0, h->status, st, ra); }));
        R_(timed(1, [&] { return gemm_bf16_tc(h->hbuf, d, l.w_o, d, M, d, d, l.b_o, h->xres, d, EPI_BIAS_RES_F32, 0, h->status, st); }));
        R_(timed(2, [&] { return launch_layernorm(h->xres, 1, l.ln2_g, l.ln2_b, h->hbuf, M, d, c.ln_eps, st); }));
        R_(timed(1, [&] { return gemm_bf16_tc(h->hbuf, d, l.w_fc1, d, M, mlp, d, l.b_fc1, h->mlp, mlp, EPI_BIAS_ACT_BF16, c.activation, h->status, st); }));
        R_(timed(1, [&] { return gemm_bf16_tc(h->mlp, mlp, l.w_fc2, mlp, M, d, mlp, l.b_fc2, h->xres, d, EPI_BIAS_RES_F32, 0, h->status, st, rf); }));
    }
    R_(timed(2, [&] { return launch_layernorm(h->xres, S, h->post_g, h->post_b, h->cls_ln, B, d, c.ln_eps, st); }));
    R_(timed(1, [&] { return gemm_bf16_tc(h->cls_ln, d, h->w_proj, d, B, D, d, nullptr, h->emb_raw, D, EPI_F32, 0, h->status, st); }));
    R_(timed(2, [&] { return launch_l2_normalize(h->emb_raw, emb_out, B, D, st); }));
#undef R_
    h->launches += 5 + 7 * (int64_t)h->layers.size();
    return B200_OK;
}

static int require_encoder(b200clip_handle* h) {
    B200_REQUIRE(h->cfg.layers > 0, "handle has no encoder (layers == 0)");
    if (h->loaded.size() != 8 + 16 * (size_t)h->cfg.layers)
        return fail(B200CLIP_ESTATE, "encoder weights are not fully loaded (call b200clip_load_weight for every tensor)");
    return B200_OK;
}

extern "C" int b200clip_encode(b200clip_handle* h, const void* patches_dev, int32_t B, float* emb_dev, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    std::lock_guard<std::mutex> lk(h->mu);
    if ((rc = require_encoder(h))) return rc;
    B200_REQUIRE(B >= 0 && (B == 0 || (patches_dev && emb_dev)), "encode: null buffer");
    const __nv_bfloat16* p = reinterpret_cast<const __nv_bfloat16*>(patches_dev);
    if ((rc = h->ws_begin((cudaStream_t)stream))) return rc;
    for (int b0 = 0; b0 < B; b0 += h->cfg.max_batch) {
        const int nb = std::min(h->cfg.max_batch, B - b0);
        if ((rc = encode_chunk(h, p + (size_t)b0 * h->n * h->kpad, nb, emb_dev + (size_t)b0 * h->cfg.embed_dim, (cudaStream_t)stream))) return rc;
    }
    return h->ws_end((cudaStream_t)stream);
}

static int embed_u8_locked(b200clip_handle* h, const uint8_t* src_dev, const int64_t* offsets, const int32_t* hw, int B, int mode,
                           float* emb_dev, cudaStream_t st) {
    int rc;
    if ((rc = require_encoder(h))) return rc;
    B200_REQUIRE((reinterpret_cast<uintptr_t>(src_dev) & 15) == 0, "embed: src_dev must be 16-byte aligned");
    if ((rc = h->ws_begin(st))) return rc;
    for (int b0 = 0; b0 < B; b0 += h->cfg.max_batch) {
        const int nb = std::min(h->cfg.max_batch, B - b0);
        cudaEvent_t pa = nullptr, pb = nullptr;
        if (h->prof) { pa = h->prof_event(); pb = h->prof_event(); cudaEventRecord(pa, st); }
        rc = preprocess_run(h->pre, src_dev, reinterpret_cast<const long long*>(offsets) + b0, hw + 2 * b0, nb, h->cfg.image_size, mode,
                            /*layout*/ 1, /*bf16*/ 1, h->cfg.patch, h->kpad, h->patches, st);
        if (h->prof) { cudaEventRecord(pb, st); h->prof_recs.push_back({0, pa, pb}); }
        if (rc) return rc;
        h->launches += 1;
        if ((rc = encode_chunk(h, h->patches, nb, emb_dev + (size_t)b0 * h->cfg.embed_dim, st))) return rc;
    }
    return h->ws_end(st);
}

extern "C" int b200clip_embed_u8(b200clip_handle* h, const uint8_t* src_dev, const int64_t* offsets, const int32_t* hw, int32_t B,
                                 int32_t mode, float* emb_dev, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    std::lock_guard<std::mutex> lk(h->mu);
    B200_REQUIRE(B >= 0 && (B == 0 || (src_dev && offsets && hw && emb_dev)), "embed_u8: null buffer");
    return embed_u8_locked(h, src_dev, offsets, hw, B, mode, emb_dev, (cudaStream_t)stream);
}

extern "C" int b200clip_embed_host(b200clip_handle* h, const uint8_t* src_host, int64_t total_bytes, const int64_t* offsets,
                                   const int32_t* hw, int32_t B, int32_t mode, float* emb_host, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    std::lock_guard<std::mutex> lk(h->mu);
    B200_REQUIRE(B >= 0 && total_bytes >= 0, "embed_host: negative size");
    if (B == 0) return B200_OK;
    B200_REQUIRE(src_host && offsets && hw && emb_host, "embed_host: null buffer");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t need = (size_t)total_bytes + 64;
    if (need > h->img_cap) {
        B200_CUDA_OK(cudaStreamSynchronize(st));
        cudaFree(h->img_dev); h->img_dev = nullptr; h->img_cap = 0;
        B200_CUDA_OK(cudaMalloc(&h->img_dev, need + need / 4));
        h->img_cap = need + need / 4;
    }
    const size_t eneed = (size_t)B * h->cfg.embed_dim;
    if (eneed > h->emb_cap) {
        B200_CUDA_OK(cudaStreamSynchronize(st));
        cudaFree(h->emb_dev); h->emb_dev = nullptr; h->emb_cap = 0;
        B200_CUDA_OK(cudaMalloc(&h->emb_dev, eneed * sizeof(float)));
        h->emb_cap = eneed;
    }
    if ((rc = h->ws_begin(st))) return rc;                 // img_dev / emb_dev are part of the shared workspace too
    B200_CUDA_OK(cudaMemcpyAsync(h->img_dev, src_host, (size_t)total_bytes, cudaMemcpyHostToDevice, st));
    if ((rc = embed_u8_locked(h, h->img_dev, offsets, hw, B, mode, h->emb_dev, st))) return rc;
    B200_CUDA_OK(cudaMemcpyAsync(emb_host, h->emb_dev, eneed * sizeof(float), cudaMemcpyDeviceToHost, st));
    B200_CUDA_OK(cudaStreamSynchronize(st));
    DeviceStatus ds;
    B200_CUDA_OK(cudaMemcpy(&ds, h->status, sizeof(ds), cudaMemcpyDeviceToHost));
    if (ds.code != 0) {
        cudaMemset(h->status, 0, sizeof(DeviceStatus));
        return fail(B200CLIP_EKERNEL, "device watchdog: pipeline barrier never completed (site " + std::to_string(ds.where) + ")");
    }
    return B200_OK;
}

// ------------------------------------------------------------------------------------------ classify / search
extern "C" int b200clip_set_labels(b200clip_handle* h, int32_t slot, const float* host_labels, int32_t N, int32_t D, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE((slot == 0 || slot == 1) && host_labels && N > 0 && D > 0, "set_labels: bad argument");
    std::lock_guard<std::mutex> lk(h->mu);
    cudaStream_t st = (cudaStream_t)stream;
    B200_CUDA_OK(cudaStreamSynchronize(st));
    cudaFree(h->labels[slot]); h->labels[slot] = nullptr;
    B200_CUDA_OK(cudaMalloc(&h->labels[slot], (size_t)N * D * sizeof(float)));
    B200_CUDA_OK(cudaMemcpyAsync(h->labels[slot], host_labels, (size_t)N * D * sizeof(float), cudaMemcpyHostToDevice, st));
    if ((rc = launch_l2_normalize(h->labels[slot], h->labels[slot], N, D, st))) return rc;
    B200_CUDA_OK(cudaStreamSynchronize(st));
    h->labels_n[slot] = N; h->labels_d[slot] = D;
    h->launches += 1;
    return B200_OK;
}

extern "C" int b200clip_gallery_pack(b200clip_handle* h, const float* src_dev, int64_t rows, int32_t D, int32_t normalize, void* dst_f16_dev, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(rows >= 0 && D > 0 && (rows == 0 || (src_dev && dst_f16_dev)), "gallery_pack: bad argument");
    rc = launch_gallery_pack(src_dev, dst_f16_dev, rows, D, normalize, (cudaStream_t)stream);
    if (rc == B200_OK && rows > 0) h->launches += 1;
    return rc;
}

extern "C" int b200clip_classify(b200clip_handle* h, int32_t slot, const float* emb_dev, int32_t B, float scale, int32_t softmax,
                                 int32_t k, float* out_values_dev, int64_t* out_index_dev, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(slot == 0 || slot == 1, "classify: slot must be 0 (labels) or 1 (scenes)");
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->labels[slot]) return fail(B200CLIP_ESTATE, "classify: no label embeddings loaded in this slot");
    rc = classify_run(emb_dev, h->labels[slot], B, h->labels_n[slot], h->labels_d[slot], scale, softmax, k, out_values_dev,
                      reinterpret_cast<long long*>(out_index_dev), (cudaStream_t)stream);
    if (rc == B200_OK && B > 0) h->launches += 1;
    return rc;
}

extern "C" int b200clip_topk(b200clip_handle* h, const void* queries, int32_t Q, const void* gallery, int64_t N, int32_t D, int32_t k,
                             int64_t index_base, float* out_scores, int64_t* out_index, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(queries && gallery && out_scores && out_index, "topk: null buffer");
    std::lock_guard<std::mutex> lk(h->mu);
    TopkPlan plan;
    if ((rc = topk_plan(Q, N, k, &plan))) return rc;
    const size_t need = plan.cand_bytes + plan.part_bytes + plan.qstage_bytes + 512;
    cudaStream_t st = (cudaStream_t)stream;
    if (need > h->topk_cap) {
        B200_CUDA_OK(cudaStreamSynchronize(st));
        cudaFree(h->topk_scratch); h->topk_scratch = nullptr; h->topk_cap = 0;
        B200_CUDA_OK(cudaMalloc(&h->topk_scratch, need));
        h->topk_cap = need;
    }
    if ((rc = h->ws_begin(st))) return rc;
    rc = topk_run(queries, gallery, Q, N, D, k, index_base, h->topk_scratch, plan, out_scores, reinterpret_cast<long long*>(out_index),
                  h->status, st);
    if (rc == B200_OK) h->launches += plan.qstage_bytes ? 3 : 2;
    if (rc == B200_OK) rc = h->ws_end(st);
    return rc;
}

extern "C" int b200clip_topk_keys(b200clip_handle* h, const void* queries, int32_t Q, const void* gallery, int64_t N, int32_t D, int32_t k,
                                  uint64_t* out_keys, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(queries && gallery && out_keys, "topk_keys: null buffer");
    std::lock_guard<std::mutex> lk(h->mu);
    TopkPlan plan;
    if ((rc = topk_plan(Q, N, k, &plan))) return rc;
    const size_t need = plan.cand_bytes + plan.part_bytes + plan.qstage_bytes + 512;
    cudaStream_t st = (cudaStream_t)stream;
    if (need > h->topk_cap) {
        B200_CUDA_OK(cudaStreamSynchronize(st));
        cudaFree(h->topk_scratch); h->topk_scratch = nullptr; h->topk_cap = 0;
        B200_CUDA_OK(cudaMalloc(&h->topk_scratch, need));
        h->topk_cap = need;
    }
    if ((rc = h->ws_begin(st))) return rc;
    rc = topk_run(queries, gallery, Q, N, D, k, 0, h->topk_scratch, plan, nullptr, nullptr, h->status, st, reinterpret_cast<unsigned long long*>(out_keys));
    if (rc == B200_OK) h->launches += plan.qstage_bytes ? 3 : 2;
    if (rc == B200_OK) rc = h->ws_end(st);
    return rc;
}

extern "C" int b200clip_merge_topk_keys(b200clip_handle* h, const uint64_t* part_keys, const int64_t* part_bases_dev, int32_t parts, int32_t Q,
                                        int32_t k, float* out_scores, int64_t* out_index, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(part_keys && part_bases_dev && out_scores && out_index, "merge_topk_keys: null buffer");
    rc = merge_topk_keys_run(reinterpret_cast<const unsigned long long*>(part_keys), reinterpret_cast<const long long*>(part_bases_dev), parts, Q, k,
                             out_scores, reinterpret_cast<long long*>(out_index), (cudaStream_t)stream);
    if (rc == B200_OK && Q > 0) h->launches += 1;
    return rc;
}

// NCCL is looked up at run time (the library does not link against it): the libnccl the process has already loaded - the one the
// caller's communicator belongs to - or, failing that, the system's.
namespace {
typedef int (*NcclAllGatherFn)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*NcclGroupFn)();
typedef const char* (*NcclErrorStringFn)(int);
struct NcclApi { NcclAllGatherFn all_gather = nullptr; NcclGroupFn group_start = nullptr, group_end = nullptr; NcclErrorStringFn error_string = nullptr; };
NcclApi g_nccl;
std::once_flag g_nccl_once;
const NcclApi* nccl_api() {
    std::call_once(g_nccl_once, [] {
        void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW);
        if (!lib) return;
        g_nccl.all_gather = reinterpret_cast<NcclAllGatherFn>(dlsym(lib, "ncclAllGather"));
        g_nccl.group_start = reinterpret_cast<NcclGroupFn>(dlsym(lib, "ncclGroupStart"));
        g_nccl.group_end = reinterpret_cast<NcclGroupFn>(dlsym(lib, "ncclGroupEnd"));
        g_nccl.error_string = reinterpret_cast<NcclErrorStringFn>(dlsym(lib, "ncclGetErrorString"));
    });
    return g_nccl.all_gather && g_nccl.group_start && g_nccl.group_end ? &g_nccl : nullptr;
}
}  // namespace

extern "C" int b200clip_search_sharded(b200clip_handle* h, const void* queries, int32_t Q, const void* gallery_shard, int64_t N_local, int32_t D,
                                       int32_t k, int64_t index_base, void* nccl_comm, int32_t world, float* out_scores, int64_t* out_index,
                                       void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(queries && out_scores && out_index && nccl_comm, "search_sharded: null buffer or communicator");
    B200_REQUIRE(Q > 0 && k > 0 && world >= 1 && N_local >= 0 && (N_local == 0 || gallery_shard), "search_sharded: bad sizes");
    const NcclApi* nccl = nccl_api();
    if (!nccl) return fail(B200CLIP_ESTATE, "search_sharded: no NCCL library in this process (libnccl.so.2 not found)");
    std::lock_guard<std::mutex> lk(h->mu);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t nk = (size_t)Q * k;
    const size_t bases_off = ((nk * 8 * ((size_t)world + 1) + 255) & ~(size_t)255);
    const size_t need_sh = bases_off + 256 + (size_t)world * 8;
    if (need_sh > h->shard_cap) {
        B200_CUDA_OK(cudaStreamSynchronize(st));
        cudaFree(h->shard_buf); h->shard_buf = nullptr; h->shard_cap = 0;
        B200_CUDA_OK(cudaMalloc(&h->shard_buf, need_sh));
        h->shard_cap = need_sh;
    }
    uint8_t* sb = static_cast<uint8_t*>(h->shard_buf);
    unsigned long long* own = reinterpret_cast<unsigned long long*>(sb);
    unsigned long long* all = own + nk;
    long long* own_base = reinterpret_cast<long long*>(sb + bases_off);
    long long* all_bases = reinterpret_cast<long long*>(sb + bases_off + 256);
    if ((rc = h->ws_begin(st))) return rc;
    if (N_local > 0) {
        TopkPlan plan;
        if ((rc = topk_plan(Q, N_local, k, &plan))) return rc;
        const size_t need = plan.cand_bytes + plan.part_bytes + plan.qstage_bytes + 512;
        if (need > h->topk_cap) {
            B200_CUDA_OK(cudaStreamSynchronize(st));
            cudaFree(h->topk_scratch); h->topk_scratch = nullptr; h->topk_cap = 0;
            B200_CUDA_OK(cudaMalloc(&h->topk_scratch, need));
            h->topk_cap = need;
        }
        if ((rc = topk_run(queries, gallery_shard, Q, N_local, D, k, 0, h->topk_scratch, plan, nullptr, nullptr, h->status, st, own))) return rc;
        h->launches += plan.qstage_bytes ? 3 : 2;
    } else {
        B200_CUDA_OK(cudaMemsetAsync(own, 0, nk * 8, st));          // an empty shard contributes empty slots
    }
    B200_CUDA_OK(cudaMemcpyAsync(own_base, &index_base, 8, cudaMemcpyHostToDevice, st));   // (pageable source: copied before the call returns)
    auto nccl_ok = [&](int r, const char* what) -> int {
        if (r == 0) return B200_OK;
        return fail(B200CLIP_ECUDA, std::string("search_sharded: ") + what + ": " + (nccl->error_string ? nccl->error_string(r) : "NCCL error"));
    };
    if ((rc = nccl_ok(nccl->group_start(), "ncclGroupStart"))) return rc;
    int r1 = nccl->all_gather(own, all, nk, /*ncclUint64*/ 5, nccl_comm, st);
    int r2 = nccl->all_gather(own_base, all_bases, 1, /*ncclInt64*/ 4, nccl_comm, st);
    int r3 = nccl->group_end();
    if ((rc = nccl_ok(r1, "ncclAllGather(keys)")) || (rc = nccl_ok(r2, "ncclAllGather(bases)")) || (rc = nccl_ok(r3, "ncclGroupEnd"))) return rc;
    rc = merge_topk_keys_run(all, all_bases, world, Q, k, out_scores, reinterpret_cast<long long*>(out_index), st);
    if (rc == B200_OK) h->launches += 1;
    if (rc == B200_OK) rc = h->ws_end(st);
    return rc;
}

extern "C" int b200clip_merge_topk(b200clip_handle* h, const float* part_scores, const int64_t* part_index, int32_t parts, int32_t Q,
                                   int32_t k, float* out_scores, int64_t* out_index, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(part_scores && part_index && out_scores && out_index, "merge_topk: null buffer");
    rc = merge_topk_run(part_scores, reinterpret_cast<const long long*>(part_index), parts, Q, k, out_scores,
                        reinterpret_cast<long long*>(out_index), (cudaStream_t)stream);
    if (rc == B200_OK && Q > 0) h->launches += 1;
    return rc;
}

// ------------------------------------------------------------------------------------------ N4: text tower
extern "C" int b200clip_text_init(b200clip_handle* h, const b200clip_text_config* cfg) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(cfg != nullptr, "text_init: null config");
    std::lock_guard<std::mutex> lk(h->mu);
    B200_REQUIRE(h->text == nullptr, "text_init: the handle already has a text tower");
    const int d = cfg->width, mlp = cfg->mlp_dim, D = cfg->embed_dim, ctx = cfg->context_length, V = cfg->vocab_size;
    B200_REQUIRE(ctx > 64 && ctx <= 4096, "text_init: context_length must be in (64, 4096] (the causal mask lives in the tcgen05 attention kernel)");
    B200_REQUIRE(d % 128 == 0 && d <= 1024 && d == cfg->heads * 64, "text_init: width must be heads*64, a multiple of 128, at most 1024");
    B200_REQUIRE(mlp % 64 == 0 && D > 0 && V > 0 && cfg->layers > 0 && cfg->max_batch > 0, "text_init: bad mlp_dim / embed_dim / vocab / layers / max_batch");
    auto* t = new TextTower();
    t->cfg = *cfg;
    t->layers.resize(cfg->layers);
    auto alloc = [&](auto** p, size_t count, bool zero) -> int {
        void* q = nullptr;
        B200_CUDA_OK(cudaMalloc(&q, count * sizeof(**p) + 256));
        if (zero) B200_CUDA_OK(cudaMemset(q, 0, count * sizeof(**p) + 256));
        t->allocs.push_back(q);
        *p = reinterpret_cast<std::remove_reference_t<decltype(*p)>>(q);
        return B200_OK;
    };
    auto bail = [&](int code) { for (void* p : t->allocs) cudaFree(p); delete t; return code; };
#define A_(call) if ((rc = (call))) return bail(rc)
    A_(alloc(&t->tok, (size_t)V * d, false)); A_(alloc(&t->pos, (size_t)ctx * d, false));
    A_(alloc(&t->final_g, d, false)); A_(alloc(&t->final_b, d, false)); A_(alloc(&t->w_proj, (size_t)D * d, false));
    for (auto& l : t->layers) {
        A_(alloc(&l.w_qkv, (size_t)3 * d * d, false)); A_(alloc(&l.b_qkv, 3 * d, false));
        A_(alloc(&l.w_o, (size_t)d * d, false)); A_(alloc(&l.b_o, d, false));
        A_(alloc(&l.w_fc1, (size_t)mlp * d, false)); A_(alloc(&l.b_fc1, mlp, false));
        A_(alloc(&l.w_fc2, (size_t)d * mlp, false)); A_(alloc(&l.b_fc2, d, false));
        A_(alloc(&l.ln1_g, d, false)); A_(alloc(&l.ln1_b, d, false)); A_(alloc(&l.ln2_g, d, false)); A_(alloc(&l.ln2_b, d, false));
    }
    const size_t Bm = cfg->max_batch, M = Bm * ctx;
    A_(alloc(&t->xres, M * d, false)); A_(alloc(&t->hbuf, M * d, true)); A_(alloc(&t->qkv, M * 3 * d, true)); A_(alloc(&t->mlp, M * mlp, true));
    A_(alloc(&t->eot, Bm * d, false)); A_(alloc(&t->eot_ln, Bm * d, true)); A_(alloc(&t->emb_raw, Bm * D, false));
#undef A_
    h->text = t;
    return B200_OK;
}

static int encode_text_chunk(b200clip_handle* h, const int32_t* ids_dev, int B, float* emb_out, cudaStream_t st) {
    TextTower* t = h->text;
    const b200clip_text_config& c = t->cfg;
    const int d = c.width, S = c.context_length, M = B * S, mlp = c.mlp_dim, D = c.embed_dim;
    int rc;
    PdlScope pdl(M <= 8192);
#define R_(call) if ((rc = (call))) return rc
    R_(launch_text_embed(ids_dev, t->tok, t->pos, t->xres, M, S, d, c.vocab_size, st));
    for (const Layer& l : t->layers) {
        R_(launch_layernorm(t->xres, 1, l.ln1_g, l.ln1_b, t->hbuf, M, d, c.ln_eps, st));
        R_(gemm_bf16_tc(t->hbuf, d, l.w_qkv, d, M, 3 * d, d, l.b_qkv, t->qkv, 3 * d, EPI_BIAS_BF16, 0, h->status, st));
        R_(launch_attention_tc(t->qkv, t->hbuf, B, S, d, c.heads, /*causal*/ 1, h->status, st));
        R_(gemm_bf16_tc(t->hbuf, d, l.w_o, d, M, d, d, l.b_o, t->xres, d, EPI_BIAS_RES_F32, 0, h->status, st));
        R_(launch_layernorm(t->xres, 1, l.ln2_g, l.ln2_b, t->hbuf, M, d, c.ln_eps, st));
        R_(gemm_bf16_tc(t->hbuf, d, l.w_fc1, d, M, mlp, d, l.b_fc1, t->mlp, mlp, EPI_BIAS_ACT_BF16, c.activation, h->status, st));
        R_(gemm_bf16_tc(t->mlp, mlp, l.w_fc2, mlp, M, d, mlp, l.b_fc2, t->xres, d, EPI_BIAS_RES_F32, 0, h->status, st));
    }
    R_(launch_text_eot_gather(ids_dev, t->xres, t->eot, B, S, d, st));
    R_(launch_layernorm(t->eot, 1, t->final_g, t->final_b, t->eot_ln, B, d, c.ln_eps, st));
    R_(gemm_bf16_tc(t->eot_ln, d, t->w_proj, d, B, D, d, nullptr, t->emb_raw, D, EPI_F32, 0, h->status, st));
    R_(launch_l2_normalize(t->emb_raw, emb_out, B, D, st));
#undef R_
    h->launches += 5 + 7 * (int64_t)t->layers.size();
    return B200_OK;
}

extern "C" int b200clip_encode_text(b200clip_handle* h, const int32_t* ids_dev, int32_t B, float* emb_dev, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    std::lock_guard<std::mutex> lk(h->mu);
    B200_REQUIRE(h->text != nullptr, "encode_text: b200clip_text_init was not called");
    if (h->text->loaded.size() != h->text->want_tensors())
        return fail(B200CLIP_ESTATE, "encode_text: text tower weights are not fully loaded");
    B200_REQUIRE(B >= 0 && (B == 0 || (ids_dev && emb_dev)), "encode_text: null buffer");
    const int mb = h->text->cfg.max_batch, ctx = h->text->cfg.context_length;
    if ((rc = h->ws_begin((cudaStream_t)stream))) return rc;
    for (int b0 = 0; b0 < B; b0 += mb) {
        const int nb = std::min(mb, B - b0);
        if ((rc = encode_text_chunk(h, ids_dev + (size_t)b0 * ctx, nb, emb_dev + (size_t)b0 * h->text->cfg.embed_dim, (cudaStream_t)stream))) return rc;
    }
    return h->ws_end((cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------ N2: JPEG decode in front of K1
extern "C" int b200clip_jpeg_probe(const void* jpeg, size_t len, int32_t* hw, int32_t* ncomp) {
    B200_REQUIRE(jpeg && hw && ncomp, "jpeg_probe: null argument");
    return jpeg_probe(jpeg, len, hw, ncomp);
}
extern "C" int b200clip_decode_jpeg(b200clip_handle* h, const void* const* jpegs, const size_t* lens, int32_t n, uint8_t* dst_dev,
                                    int64_t dst_capacity, int64_t* offsets_out, int32_t* hw_out, int32_t* status_out, int64_t* total_out,
                                    int32_t host_threads, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(n >= 0 && (n == 0 || (jpegs && lens && offsets_out && hw_out && status_out)) && total_out && dst_capacity >= 0, "decode_jpeg: bad argument");
    std::lock_guard<std::mutex> lock(h->mu);
    if (!h->jpeg) { h->jpeg = jpeg_state_create(&rc); if (!h->jpeg) return rc; }
    int launches = 0;
    if ((rc = h->ws_begin((cudaStream_t)stream))) return rc;
    rc = jpeg_decode_batch(h->jpeg, jpegs, lens, n, dst_dev, (size_t)dst_capacity, offsets_out, hw_out, status_out, total_out, host_threads,
                           (cudaStream_t)stream, &launches);
    h->launches += launches;
    if (rc == B200_OK) rc = h->ws_end((cudaStream_t)stream);
    return rc;
}

// ------------------------------------------------------------------------------------------ building blocks (tests)
extern "C" int b200clip_gemm_bf16(b200clip_handle* h, const void* A, const void* W, int32_t M, int32_t N, int32_t K, const float* bias,
                                  void* out, int32_t epilogue, int32_t activation, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(A && W && out && epilogue >= 0 && epilogue <= 3, "gemm: bad argument");
    rc = gemm_bf16_tc(A, K, W, K, M, N, K, bias, out, N, epilogue, activation, h->status, (cudaStream_t)stream);
    if (rc == B200_OK) h->launches += 1;
    return rc;
}
extern "C" int b200clip_attention(b200clip_handle* h, const void* qkv, void* out, int32_t B, int32_t S, int32_t width, int32_t heads, int32_t causal,
                                  void* stream) {
    int rc = set_device(h); if (rc) return rc;
    B200_REQUIRE(!causal || (S > 64 && width == heads * 64), "attention: the causal mask needs S > 64 and head_dim 64 (tcgen05 kernel)");
    rc = launch_attention_auto(qkv, out, B, S, width, heads, causal, h->status, (cudaStream_t)stream);
    if (rc == B200_OK) h->launches += 1;
    return rc;
}
extern "C" int b200clip_layernorm(b200clip_handle* h, const float* x, const float* g, const float* b, void* out, int32_t rows, int32_t width,
                                  float eps, void* stream) {
    int rc = set_device(h); if (rc) return rc;
    rc = launch_layernorm(x, 1, g, b, out, rows, width, eps, (cudaStream_t)stream);
    if (rc == B200_OK) h->launches += 1;
    return rc;
}
