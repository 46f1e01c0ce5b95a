// view_graph.cu — CUDA-graph replay of ashcuda_view_run ("CUDA-graph replay for the whole-view number").
//
// One view is ~19 small kernels on two streams; issuing them costs the host ~110 us against ~235 us of GPU
// time (tools/issue_cost.py), so a host thread that also feeds uploads is close to being the bottleneck.
// The whole view is therefore captured ONCE per TAA ping-pong parity (history / output swap every view)
// and replayed with a single cudaGraphLaunch.  What changes from view to view are a few by-value kernel
// arguments — the cull matrix, the TAA jitter / blend constants, the LOD camera, the camera block and the
// view id — and exactly those kernel nodes are patched before each replay with
// cudaGraphExecKernelNodeSetParams (the kernels are the same ones the eager path launches: nothing in
// them knows about graphs).  Everything else a captured view depends on (buffers, counts, stage mask,
// bloom / tonemap constants, the TAA kernel variant) is in GraphKey; a mismatch re-captures.
//
// The camera block and the view id used to be small host->device copies from pageable memory, which a
// graph would freeze at capture time; both are now written by k_store_view from by-value arguments.
#include <cstring>
#include <string>
#include <vector>
#include "ash_internal.cuh"

namespace ash {

__global__ void k_store_view(AshCullingCameraData cam, uint32_t with_camera, uint32_t view_id,
                             AshCullingCameraData* __restrict__ camera_dev, AshViewStats* __restrict__ stats) {
    if (with_camera) {
        const float* src = reinterpret_cast<const float*>(&cam);
        float* dst = reinterpret_cast<float*>(camera_dev);
        for (uint32_t i = threadIdx.x; i < sizeof(AshCullingCameraData) / 4; i += blockDim.x) dst[i] = src[i];
    }
    if (threadIdx.x == 0) stats->view_id = view_id;
}

cudaError_t launch_store_view(const AshCullingCameraData& cam, bool with_camera, uint32_t view_id,
                              AshCullingCameraData* camera_dev, AshViewStats* stats, const Launch& L) {
    k_store_view<<<1, 64, 0, L.s>>>(cam, with_camera ? 1u : 0u, view_id, camera_dev, stats);
    *L.launches += 1;
    return cudaGetLastError();
}

namespace {

constexpr int kMaxArgs = 16;

GraphKey make_key(const ashcuda_ctx* c, const AshCudaViewParams* vp) {
    GraphKey k;
    std::memset(&k, 0, sizeof(k));                       // padding too: the key is compared with memcmp
    k.stream = c->stream;
    k.depth = c->depth; k.color = c->color; k.velocity = c->velocity;
    k.objs = c->objs; k.draws_in = c->draws_in; k.lights = c->lights;
    k.inst = c->inst; k.inst_batch = c->inst_batch; k.inst_mesh = c->inst_mesh; k.lod_meshes = c->lod_meshes;
    k.stages = vp->stages;
    k.n_objs = c->n_objs; k.n_draws = c->n_draws; k.n_lights = c->n_lights; k.n_inst = c->n_inst; k.n_lod_meshes = c->n_lod_meshes;
    k.object_count = vp->cull.object_count; k.hiz_levels = vp->cull.hiz_levels;
    k.lod_use_visibility = c->lod_cfg.use_visibility;
    k.light = vp->light;
    std::memcpy(k.taa_screen, vp->taa.screen_params, sizeof(k.taa_screen));
    k.bloom = vp->bloom; k.post = vp->post;
    k.in_halo = taa_in_halo(vp->taa) ? 1u : 0u;
    k.tonemap_op = c->tonemap_op;
    k.fwd_velocity = c->fwd_has_reproj ? 1u : 0u;
    if (vp->stages & ASH_STAGE_LIGHTS) {
        k.planes_cached = (c->planes_valid && std::memcmp(c->planes_ip, vp->camera.inv_projection, sizeof(c->planes_ip)) == 0) ? 1u : 0u;
        std::memcpy(k.inv_projection, vp->camera.inv_projection, sizeof(k.inv_projection));
    }
    if (c->cfg.precision == ASH_PRECISION_FAST && (vp->stages & ASH_STAGE_TAA) && k.in_halo) {   // which k_taa_tol instance the graph holds
        TaaTolArgs ta;
        k.tol_variant = 1u + ash_taa_tol_plan(const_cast<ashcuda_ctx*>(c), vp->taa, vp->bloom.threshold, &ta).regular;
    }
    k.taa_cur = (uint32_t)c->taa_cur; k.history_valid = c->history_valid; k.taa_ran = c->taa_ran;
    k.have_color = c->have_color; k.have_depth = c->have_depth; k.have_instancing = c->have_instancing;
    k.concurrent = c->cfg.concurrent_frontend; k.have_pyramid = c->have_pyramid;
    k.fwd = c->fwd_params;
    return k;
}

// the per-view arguments of this view, in the order ViewGraph keeps the nodes
struct Patches { PatchArg store, cull, cull2, taa, lod, fwd, vel; bool has_cull, has_cull2, has_taa, has_lod, has_fwd, has_vel; };

void make_patches(const ashcuda_ctx* c, const AshCudaViewParams* vp, Patches* P) {
    const uint32_t st = vp->stages;
    P->has_cull = (st & ASH_STAGE_CULL) && vp->cull.object_count > 0;
    P->has_taa = (st & ASH_STAGE_TAA) != 0;
    P->has_lod = (st & ASH_STAGE_INSTANCING) && vp->cull.object_count > 0;
    P->has_cull2 = P->has_cull && (st & ASH_STAGE_TWO_PHASE);
    if (P->has_cull) cull_patch(vp->cull, P->has_cull2 ? 1 : 0, &P->cull);
    if (P->has_cull2) cull_patch(vp->cull, 2, &P->cull2);
    if (P->has_taa) {
        const bool vel = (st & ASH_STAGE_FORWARD_SHADE) ? c->fwd_has_reproj : c->velocity != nullptr;
        if (c->cfg.precision == ASH_PRECISION_FAST && !vel && taa_in_halo(vp->taa)) {
            TaaTolArgs ta;
            taa_tol_patch((st & ASH_STAGE_BLOOM) != 0, ash_taa_tol_plan(const_cast<ashcuda_ctx*>(c), vp->taa, vp->bloom.threshold, &ta), &P->taa);
        }
        else taa_pair_patch((st & ASH_STAGE_BLOOM) != 0, vel, vp->taa, &P->taa);
    }
    if (P->has_lod) {
        AshInstancingParams ip{};
        std::memcpy(ip.view_proj, vp->cull.view_proj, sizeof(ip.view_proj));
        std::memcpy(ip.camera_pos, vp->camera.camera_pos, sizeof(ip.camera_pos));
        ip.screen_size[0] = vp->cull.screen_params[0]; ip.screen_size[1] = vp->cull.screen_params[1];
        ip.object_count = vp->cull.object_count;
        lod_patch(ip, c->lod_cfg, &P->lod);
    }
    P->has_fwd = (st & ASH_STAGE_FORWARD_DEPTH) != 0;
    if (P->has_fwd) forward_setup_patch(vp->camera, &P->fwd);
    P->has_vel = (st & ASH_STAGE_FORWARD_SHADE) && c->fwd_has_reproj;
    if (P->has_vel) forward_velocity_patch(c->fwd_reproj, &P->vel);
    P->store.func = reinterpret_cast<const void*>(&k_store_view);
    P->store.n_args = 5;
}

cudaError_t find_node(cudaGraph_t g, const void* func, cudaGraphNode_t* out) {
    size_t n = 0;
    cudaError_t e = cudaGraphGetNodes(g, nullptr, &n);
    if (e != cudaSuccess) return e;
    std::vector<cudaGraphNode_t> nodes(n);
    e = cudaGraphGetNodes(g, nodes.data(), &n);
    if (e != cudaSuccess) return e;
    *out = nullptr;
    int found = 0;
    for (cudaGraphNode_t nd : nodes) {
        cudaGraphNodeType t;
        if (cudaGraphNodeGetType(nd, &t) != cudaSuccess || t != cudaGraphNodeTypeKernel) continue;
        cudaKernelNodeParams kp;
        if (cudaGraphKernelNodeGetParams(nd, &kp) != cudaSuccess) continue;
        if (kp.func == func) { *out = nd; ++found; }
    }
    return found == 1 ? cudaSuccess : cudaErrorInvalidValue;      // a per-view kernel must appear exactly once
}

// replace one by-value argument of a kernel node of the instantiated graph
cudaError_t patch_node(ViewGraph& g, cudaGraphNode_t node, int n_args, int idx0, void* v0, int idx1 = -1, void* v1 = nullptr,
                       int idx2 = -1, void* v2 = nullptr) {
    cudaKernelNodeParams kp;
    cudaError_t e = cudaGraphKernelNodeGetParams(node, &kp);      // argument storage owned by the (kept) source graph
    if (e != cudaSuccess) return e;
    if (n_args > kMaxArgs || !kp.kernelParams) return cudaErrorInvalidValue;
    void* args[kMaxArgs];
    for (int i = 0; i < n_args; ++i) args[i] = kp.kernelParams[i];
    args[idx0] = v0;
    if (idx1 >= 0) args[idx1] = v1;
    if (idx2 >= 0) args[idx2] = v2;
    kp.kernelParams = args;
    kp.extra = nullptr;
    return cudaGraphExecKernelNodeSetParams(g.exec, node, &kp);
}

void drop(ViewGraph& g) {
    if (g.exec) cudaGraphExecDestroy(g.exec);
    if (g.graph) cudaGraphDestroy(g.graph);
    g = ViewGraph{};
}

struct StateSnap {
    int taa_cur; bool history_valid, taa_ran, bloom_valid, have_pyramid, have_camera; uint64_t launches; AshCullingCameraData camera_host;
    float taa_chk[4]; bool taa_exact; float planes_ip[4][4]; bool planes_valid;
    bool have_depth, have_color, have_forward; const float* depth; const void* color; const void* velocity;
};
StateSnap snap(const ashcuda_ctx* c) {
    StateSnap s{};
    s.taa_cur = c->taa_cur; s.history_valid = c->history_valid; s.taa_ran = c->taa_ran; s.bloom_valid = c->bloom_valid;
    s.have_pyramid = c->have_pyramid; s.have_camera = c->have_camera; s.launches = c->launches; s.camera_host = c->camera_host; s.taa_exact = c->taa_exact;
    std::memcpy(s.taa_chk, c->taa_chk, sizeof(s.taa_chk));
    s.have_depth = c->have_depth; s.have_color = c->have_color; s.have_forward = c->have_forward;
    s.depth = c->depth; s.color = c->color; s.velocity = c->velocity;
    std::memcpy(s.planes_ip, c->planes_ip, sizeof(s.planes_ip)); s.planes_valid = c->planes_valid;
    return s;
}
void restore(ashcuda_ctx* c, const StateSnap& s) {
    c->taa_cur = s.taa_cur; c->history_valid = s.history_valid; c->taa_ran = s.taa_ran; c->bloom_valid = s.bloom_valid;
    c->have_pyramid = s.have_pyramid; c->have_camera = s.have_camera; c->launches = s.launches; c->camera_host = s.camera_host;
    std::memcpy(c->taa_chk, s.taa_chk, sizeof(s.taa_chk)); c->taa_exact = s.taa_exact;
    c->have_depth = s.have_depth; c->have_color = s.have_color; c->have_forward = s.have_forward;
    c->depth = s.depth; c->pyr.mip0 = c->depth; c->color = s.color; c->velocity = s.velocity;
    std::memcpy(c->planes_ip, s.planes_ip, sizeof(c->planes_ip)); c->planes_valid = s.planes_valid;
}

}  // namespace
}  // namespace ash

using namespace ash;

void ash_view_graphs_destroy(ashcuda_ctx* c) {
    if (!c) return;
    drop(c->vgraph[0]);
    drop(c->vgraph[1]);
}

// Runs the view through a captured graph when it can.  *handled = false means "not a graph case, run
// the eager path" (nothing was launched and no state changed).
int ash_view_run_graph(ashcuda_ctx* c, const AshCudaViewParams* vp, bool* handled) {
    *handled = false;
    const uint32_t st = vp->stages;
    if (!c->cfg.graph_replay || c->graphs_broken || c->prof.mask != 0 || c->cfg.fp32_intermediates) return ASH_OK;
    // the TAA node must be the pair kernel (post_fast.cu dispatch: exact taps and (velocity || jitter delta inside the halo))
    const bool shaded = (st & ASH_STAGE_FORWARD_SHADE) != 0;                     // colour made inside the view, no velocity
    if (st & ASH_STAGE_TAA) {
        if (!(c->have_color || shaded) || !c->cfg.enable_post) return ASH_OK;    // eager path reports the error
        if (!((shaded ? c->fwd_has_reproj : c->velocity != nullptr) || taa_in_halo(vp->taa))) return ASH_OK;
    }
    ViewGraph& g = c->vgraph[c->taa_cur & 1];
    const GraphKey key = make_key(c, vp);
    Patches P;
    make_patches(c, vp, &P);

    if (!(g.valid && std::memcmp(&g.key, &key, sizeof(key)) == 0)) {
        // ---- capture: run the eager body once into a graph (nothing executes yet)
        drop(g);
        if (cudaSetDevice(c->device) != cudaSuccess) return ASH_OK;
        if ((st & ASH_STAGE_FORWARD_DEPTH) && ash_forward_alloc(c) != ASH_OK) return ASH_OK;   // lazy buffers: not inside a capture
        const StateSnap s0 = snap(c);
        if (cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); c->graphs_broken = true; return ASH_OK; }
        const int rc = ash_view_run_eager(c, vp);
        cudaGraph_t graph = nullptr;
        cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
        bool ok = rc == ASH_OK && e == cudaSuccess && graph;
        if (ok) {
            g.graph = graph;
            ok = cudaGraphInstantiate(&g.exec, g.graph, 0) == cudaSuccess;
            ok = ok && find_node(g.graph, P.store.func, &g.n_store) == cudaSuccess;
            if (ok && P.has_cull) ok = find_node(g.graph, P.cull.func, &g.n_cull) == cudaSuccess;
            if (ok && P.has_cull2) ok = find_node(g.graph, P.cull2.func, &g.n_cull2) == cudaSuccess;
            if (ok && P.has_taa) ok = find_node(g.graph, P.taa.func, &g.n_taa) == cudaSuccess;     // absent: TAA taps not exact
            if (ok && P.has_lod) ok = find_node(g.graph, P.lod.func, &g.n_lod) == cudaSuccess;
            if (ok && P.has_fwd) ok = find_node(g.graph, P.fwd.func, &g.n_fwd) == cudaSuccess;
            if (ok && P.has_vel) ok = find_node(g.graph, P.vel.func, &g.n_vel) == cudaSuccess;
        } else if (graph) cudaGraphDestroy(graph);
        if (!ok) {
            // an invalid view (rc != 0) is the caller's error, reported by the eager path below; a failed capture of a
            // valid view switches replay off for this context
            cudaGetLastError();
            drop(g);
            restore(c, s0);
            if (rc == ASH_OK) c->graphs_broken = true;
            return ASH_OK;
        }
        g.key = key; g.valid = true;
        g.launches = c->launches - s0.launches;
        g.taa_flip = c->taa_cur != s0.taa_cur;
        g.history_valid = c->history_valid; g.taa_ran = c->taa_ran; g.bloom_valid = c->bloom_valid;
        g.have_pyramid = c->have_pyramid; g.have_camera = c->have_camera;
        g.have_depth = c->have_depth; g.have_color = c->have_color; g.have_forward = c->have_forward;
        g.depth = c->depth; g.color = c->color; g.velocity = c->velocity;
        c->graph_captures += 1;
        // the capture left the context in its post-view state and the argument values of this very view in the
        // nodes: launch as captured
        cudaError_t le = cudaGraphLaunch(g.exec, c->stream);
        if (le != cudaSuccess) {
            cudaGetLastError(); drop(g); restore(c, s0); c->graphs_broken = true;
            return ASH_OK;
        }
        *handled = true;
        return ASH_OK;
    }

    // ---- replay: patch the per-view arguments, one launch
    ASH_CUDA(c, cudaSetDevice(c->device));
    AshCullingCameraData cam = vp->camera;
    uint32_t with_camera = (st & ASH_STAGE_LIGHTS) ? 1u : 0u, view_id = vp->view_id;
    ASH_CUDA(c, patch_node(g, g.n_store, P.store.n_args, 0, &cam, 1, &with_camera, 2, &view_id));
    if (P.has_cull) ASH_CUDA(c, patch_node(g, g.n_cull, P.cull.n_args, P.cull.arg_index, P.cull.value));
    if (P.has_cull2) ASH_CUDA(c, patch_node(g, g.n_cull2, P.cull2.n_args, P.cull2.arg_index, P.cull2.value));
    if (P.has_taa) ASH_CUDA(c, patch_node(g, g.n_taa, P.taa.n_args, P.taa.arg_index, P.taa.value));
    if (P.has_lod) ASH_CUDA(c, patch_node(g, g.n_lod, P.lod.n_args, P.lod.arg_index, P.lod.value));
    if (P.has_fwd) ASH_CUDA(c, patch_node(g, g.n_fwd, P.fwd.n_args, P.fwd.arg_index, P.fwd.value));
    if (P.has_vel) ASH_CUDA(c, patch_node(g, g.n_vel, P.vel.n_args, P.vel.arg_index, P.vel.value));
    ASH_CUDA(c, cudaGraphLaunch(g.exec, c->stream));
    if ((st & ASH_STAGE_LIGHTS) && vp->light.light_count != 0u) {      // the tile side planes now match this invProjection
        std::memcpy(c->planes_ip, vp->camera.inv_projection, sizeof(c->planes_ip)); c->planes_valid = true;
    }
    if (g.taa_flip) c->taa_cur ^= 1;
    c->history_valid = g.history_valid; c->taa_ran = g.taa_ran; c->bloom_valid = g.bloom_valid;
    c->have_pyramid = g.have_pyramid; c->have_camera = g.have_camera;
    c->have_depth = g.have_depth; c->have_color = g.have_color; c->have_forward = g.have_forward;
    c->depth = g.depth; c->pyr.mip0 = c->depth; c->color = g.color; c->velocity = g.velocity;
    if (st & ASH_STAGE_LIGHTS) c->camera_host = vp->camera;
    c->launches += g.launches;
    c->graph_replays += 1;
    *handled = true;
    return ASH_OK;
}
