// K7: PathManager::filterUnique (reference src/PathManager.cpp:432-438) and path export.
// std::unique drops a path when its node sequence equals the previous path's (Path::operator==,
// src/Path.cpp:46-49, edges are not compared): only CONSECUTIVE duplicates disappear.
#include "ctx.cuh"
#include "primitives.cuh"

namespace hera {

// One warp per path: keep[i] = 0 iff path i has the same node sequence as path i-1.
__global__ void __launch_bounds__(128) unique_flags_kernel(uint64_t paths, const uint64_t *__restrict__ off, const uint32_t *__restrict__ nodes,
                                                           uint8_t *__restrict__ keep, uint32_t *__restrict__ kept_nodes) {
    const uint64_t i = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = lane_id();
    if (i >= paths) return;
    const uint64_t b = off[i], n = off[i + 1] - b;
    bool same = false;
    if (i > 0) {
        const uint64_t pb = off[i - 1], pn = b - pb;
        if (pn == n) {
            bool diff = false;
            for (uint64_t k = lane; k < n && !diff; k += 32) diff = nodes[b + k] != nodes[pb + k];
            same = !__any_sync(0xffffffffu, diff);
        }
    }
    if (lane == 0) { keep[i] = same ? 0 : 1; kept_nodes[i] = same ? 0u : static_cast<uint32_t>(n); }
}

__global__ void __launch_bounds__(128) unique_copy_kernel(uint64_t paths, const uint64_t *__restrict__ off, const uint32_t *__restrict__ nodes,
                                                          const uint32_t *__restrict__ slots, const int64_t *__restrict__ len, const uint8_t *__restrict__ keep,
                                                          const uint32_t *__restrict__ path_scan, const uint64_t *__restrict__ node_scan,
                                                          uint64_t *__restrict__ u_off, uint32_t *__restrict__ u_nodes, uint32_t *__restrict__ u_slots,
                                                          int64_t *__restrict__ u_len) {
    const uint64_t i = (static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = lane_id();
    if (i >= paths || !keep[i]) return;
    const uint64_t b = off[i], n = off[i + 1] - b, p = path_scan[i], d = node_scan[i];
    if (lane == 0) { u_off[p + 1] = d + n; u_len[p] = len[i]; if (p == 0) u_off[0] = 0; }
    for (uint64_t k = lane; k < n; k += 32) u_nodes[d + k] = nodes[b + k];
    for (uint64_t k = lane; k + 1 < n; k += 32) u_slots[d - p + k] = slots[b - i + k];
}

void unique_paths(Ctx &c, uint64_t *n_unique) {
    PathStore &a = c.all, &u = c.uniq;
    u.paths = 0; u.nodes = 0;
    if (a.paths > 0) {
        TempPool tmp(c.stream);
        uint8_t *keep = tmp.get<uint8_t>(a.paths);
        uint32_t *kept_nodes = tmp.get<uint32_t>(a.paths), *path_scan = tmp.get<uint32_t>(a.paths), *d_np = tmp.get<uint32_t>(1);
        uint64_t *node_scan = tmp.get<uint64_t>(a.paths), *d_nn = tmp.get<uint64_t>(1);
        HERA_LAUNCH(c, unique_flags_kernel, div_up(a.paths * 32, 128), 128, 0, a.paths, a.off.p, a.node.p, keep, kept_nodes);
        exclusive_scan_u8_to_u32(c, keep, a.paths, path_scan, d_np);
        exclusive_scan_u32_to_u64(c, kept_nodes, a.paths, node_scan, d_nn);
        uint32_t np = 0;
        uint64_t nn = 0;
        HERA_CUDA(cudaMemcpyAsync(&np, d_np, sizeof(np), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaMemcpyAsync(&nn, d_nn, sizeof(nn), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
        u.off.reserve(np + 1); u.len.reserve(np); u.node.reserve(nn); u.slot.reserve(nn);
        HERA_LAUNCH(c, unique_copy_kernel, div_up(a.paths * 32, 128), 128, 0, a.paths, a.off.p, a.node.p, a.slot.p, a.len.p, keep, path_scan, node_scan,
                    u.off.p, u.node.p, u.slot.p, u.len.p);
        u.paths = np; u.nodes = nn;
        HERA_CUDA(cudaStreamSynchronize(c.stream));
    }
    c.uniq_ready = true;
    c.groups_valid = false;
    if (n_unique) *n_unique = u.paths;
}

void export_paths(Ctx &c, hera_paths_view *out) {
    const PathStore &ps = c.uniq_ready ? c.uniq : c.all;
    auto &m = c.pm;
    *out = hera_paths_view{};
    m.off.reserve(ps.paths + 1); m.len.reserve(ps.paths + 1); m.nodes.reserve(ps.nodes + 1); m.slots.reserve(ps.nodes + 1);
    m.off.p[0] = 0;
    if (ps.paths) {
        HERA_CUDA(cudaMemcpyAsync(m.off.p, ps.off.p, (ps.paths + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaMemcpyAsync(m.len.p, ps.len.p, ps.paths * sizeof(int64_t), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaMemcpyAsync(m.nodes.p, ps.node.p, ps.nodes * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
        if (ps.nodes > ps.paths)
            HERA_CUDA(cudaMemcpyAsync(m.slots.p, ps.slot.p, (ps.nodes - ps.paths) * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream));
        HERA_CUDA(cudaStreamSynchronize(c.stream));
    }
    out->paths = ps.paths; out->total_nodes = ps.nodes;
    out->off = m.off.p; out->nodes = m.nodes.p; out->slots = m.slots.p; out->len = m.len.p;
    out->calls = c.call_kind.size();
    out->call_kind = c.call_kind.data(); out->call_metric = c.call_metric.data(); out->call_found = c.call_found.data();
    out->mc_walks = c.mc_draws.size();
    out->mc_draws_per_walk = c.mc_draws.data();
}

}  // namespace hera
