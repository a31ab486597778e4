// Vertex graph colouring ON THE DEVICE: for meshes too large for an offline pass and for
// re-colouring what a partitioner or a generator produced. The host restatement of the reference tool lives in coloring.cpp
// (OrderedGreedy / Mcs / balanceColoredCategories, M/GraphColoring/*, pinned node by node); this file is the parallel
// counterpart with the same contract - a valid colouring, few colours, balanced categories - not the same colours:
//   * speculative greedy (Gebremedhin-Manne): every uncoloured node takes the smallest colour its committed neighbours do
//     not hold (64-bit forbidden mask, read from the previous round's snapshot), then conflicts between nodes coloured in the
//     same round are resolved by a fixed priority (larger degree first, lower id on ties): the loser tries again next round.
//     Everything a round reads was written by the previous kernel, so the result is deterministic;
//   * balancing follows the tool's rule (move nodes from the largest category to the smallest, M/GraphColoring/
//     ColoringAlgorithms.cpp:152-198) in batches: all nodes of the source colour without a neighbour in the target colour are
//     movable AT ONCE (they share a colour, so no two of them are adjacent); the first k by node id move, k = half the gap,
//     ranks from a device prefix sum - deterministic as well.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <mutex>
#include <vector>
#include <cuda_runtime.h>
#include <cub/cub.cuh>

#include "../../include/gaia_b200.h"

namespace {

#define CKI(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) { cleanup(); return GAIA_ERR_CUDA; } } while (0)

__device__ __forceinline__ bool beats(int du, int u, int dv, int v) { return du > dv || (du == dv && u < v); }   // u has priority over v

// tentative colour for every node that has none: smallest colour no COMMITTED neighbour holds
__global__ void k_speculate(int n, const int32_t* __restrict__ off, const int32_t* __restrict__ adj, const int32_t* __restrict__ committed,
                            int32_t* __restrict__ tentative, int32_t* overflow) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    if (committed[v] >= 0) { tentative[v] = committed[v]; return; }
    unsigned long long forbidden = 0ull;
    for (int q = off[v]; q < off[v + 1]; q++) { const int c = committed[adj[q]]; if (c >= 0) forbidden |= 1ull << c; }
    if (forbidden == ~0ull) { *overflow = 1; tentative[v] = -1; return; }
    tentative[v] = __ffsll((long long)~forbidden) - 1;
}
// a node keeps its tentative colour unless a neighbour that was ALSO uncoloured picked the same one and has priority
__global__ void k_resolve(int n, const int32_t* __restrict__ off, const int32_t* __restrict__ adj, const int32_t* __restrict__ committed,
                          const int32_t* __restrict__ tentative, int32_t* __restrict__ next, int32_t* nLeft) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    if (committed[v] >= 0) { next[v] = committed[v]; return; }
    const int c = tentative[v], dv = off[v + 1] - off[v];
    bool keep = c >= 0;
    for (int q = off[v]; keep && q < off[v + 1]; q++) {
        const int u = adj[q];
        if (committed[u] < 0 && tentative[u] == c && beats(off[u + 1] - off[u], u, dv, v)) keep = false;
    }
    next[v] = keep ? c : -1;
    if (!keep) atomicAdd(nLeft, 1);
}
__global__ void k_histogram(int n, const int32_t* __restrict__ col, int32_t* sizes) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) atomicAdd(sizes + col[v], 1);
}
// movable[v] = 1 when v has colour `from` and no neighbour of colour `to`
__global__ void k_movable(int n, const int32_t* __restrict__ off, const int32_t* __restrict__ adj, const int32_t* __restrict__ col, int from, int to,
                          int32_t* __restrict__ movable) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    int m = col[v] == from;
    for (int q = off[v]; m && q < off[v + 1]; q++) if (col[adj[q]] == to) m = 0;
    movable[v] = m;
}
__global__ void k_move(int n, const int32_t* __restrict__ movable, const int32_t* __restrict__ rank, int k, int to, int32_t* col) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n && movable[v] && rank[v] < k) col[v] = to;
}

}  // namespace

extern "C" int gaia_b200_color_graph_device(int device, int32_t numNodes, int64_t nEdges, const int32_t* edges, int32_t flags,
                                            float goalMaxMinRatio, int32_t* colors, int32_t* categoryOffsets, int32_t* categoryNodes) {
    if (numNodes < 0 || nEdges < 0 || (nEdges > 0 && !edges)) return GAIA_ERR_INVALID_ARG;
    if (numNodes == 0) return 0;
    int nDev = 0;
    if (cudaGetDeviceCount(&nDev) != cudaSuccess || device < 0 || device >= nDev) return GAIA_ERR_NO_DEVICE;
    if (cudaSetDevice(device) != cudaSuccess) return GAIA_ERR_CUDA;
    // CSR of the undirected graph (unique edges)
    std::vector<std::pair<int32_t, int32_t>> e((size_t)nEdges);
    for (int64_t i = 0; i < nEdges; i++) {
        int32_t a = edges[2 * i], b = edges[2 * i + 1];
        if (a < 0 || b < 0 || a >= numNodes || b >= numNodes || a == b) return GAIA_ERR_INVALID_ARG;
        if (a > b) std::swap(a, b);
        e[i] = {a, b};
    }
    std::sort(e.begin(), e.end());
    e.erase(std::unique(e.begin(), e.end()), e.end());
    const int n = numNodes;
    std::vector<int32_t> off(n + 1, 0), adj(2 * e.size());
    for (auto& p : e) { off[p.first + 1]++; off[p.second + 1]++; }
    for (int i = 0; i < n; i++) off[i + 1] += off[i];
    {
        std::vector<int32_t> cur(off.begin(), off.end() - 1);
        for (auto& p : e) { adj[cur[p.first]++] = p.second; adj[cur[p.second]++] = p.first; }
    }
    int32_t *dOff = nullptr, *dAdj = nullptr, *dA = nullptr, *dB = nullptr, *dT = nullptr, *dCnt = nullptr, *dSizes = nullptr, *dRank = nullptr;
    void* dTmp = nullptr;
    auto cleanup = [&] { cudaFree(dOff); cudaFree(dAdj); cudaFree(dA); cudaFree(dB); cudaFree(dT); cudaFree(dCnt); cudaFree(dSizes); cudaFree(dRank); cudaFree(dTmp); };
    CKI(cudaMalloc(&dOff, (n + 1) * sizeof(int32_t))); CKI(cudaMalloc(&dAdj, std::max<size_t>(adj.size(), 1) * sizeof(int32_t)));
    CKI(cudaMalloc(&dA, n * sizeof(int32_t))); CKI(cudaMalloc(&dB, n * sizeof(int32_t))); CKI(cudaMalloc(&dT, n * sizeof(int32_t)));
    CKI(cudaMalloc(&dRank, n * sizeof(int32_t))); CKI(cudaMalloc(&dCnt, 2 * sizeof(int32_t))); CKI(cudaMalloc(&dSizes, 64 * sizeof(int32_t)));
    CKI(cudaMemcpy(dOff, off.data(), (n + 1) * sizeof(int32_t), cudaMemcpyHostToDevice));
    if (!adj.empty()) CKI(cudaMemcpy(dAdj, adj.data(), adj.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    CKI(cudaMemset(dA, 0xff, n * sizeof(int32_t)));
    const int B = 256, G = (n + B - 1) / B;
    int32_t *cur = dA, *nxt = dB;
    for (int round = 0; round < 10000; round++) {
        int32_t h[2] = {0, 0};
        CKI(cudaMemset(dCnt, 0, 2 * sizeof(int32_t)));
        k_speculate<<<G, B>>>(n, dOff, dAdj, cur, dT, dCnt + 1);
        k_resolve<<<G, B>>>(n, dOff, dAdj, cur, dT, nxt, dCnt);
        CKI(cudaMemcpy(h, dCnt, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost));
        std::swap(cur, nxt);
        if (h[1]) { cleanup(); return GAIA_ERR_UNSUPPORTED; }      // more than 64 colours
        if (!h[0]) break;
    }
    // category sizes
    std::vector<int32_t> sizes(64, 0);
    auto histogram = [&]() -> cudaError_t {
        cudaError_t e1 = cudaMemset(dSizes, 0, 64 * sizeof(int32_t));
        if (e1 != cudaSuccess) return e1;
        k_histogram<<<G, B>>>(n, cur, dSizes);
        return cudaMemcpy(sizes.data(), dSizes, 64 * sizeof(int32_t), cudaMemcpyDeviceToHost);
    };
    CKI(histogram());
    int nc = 0;
    for (int c = 0; c < 64; c++) if (sizes[c]) nc = c + 1;
    if ((flags & GAIA_COLOR_BALANCE) && nc > 1) {
        size_t tmpBytes = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tmpBytes, dT, dRank, n);
        CKI(cudaMalloc(&dTmp, tmpBytes + 16));
        for (int round = 0; round < 256 * nc; round++) {
            std::vector<int> order(nc);
            for (int c = 0; c < nc; c++) order[c] = c;
            std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return sizes[a] > sizes[b]; });      // largest first
            if ((float)sizes[order[0]] <= goalMaxMinRatio * (float)std::max(sizes[order[nc - 1]], 1)) break;
            // the tool moves from the largest category to the smallest and falls back to other sources; in batches a target can
            // be saturated (no movable node left for it), so targets are tried smallest first and sources largest first
            bool moved = false;
            for (int ti = nc - 1; ti > 0 && !moved; ti--) {
                const int to = order[ti];
                for (int si = 0; si < ti && !moved; si++) {
                    const int from = order[si], gap = sizes[from] - sizes[to];
                    if (gap < 2) break;
                    k_movable<<<G, B>>>(n, dOff, dAdj, cur, from, to, dT);
                    CKI(cub::DeviceScan::ExclusiveSum(dTmp, tmpBytes, dT, dRank, n));
                    k_move<<<G, B>>>(n, dT, dRank, gap / 2, to, cur);
                    const int32_t before = sizes[to];
                    CKI(histogram());
                    moved = sizes[to] != before;
                }
            }
            if (!moved) break;                // "not optimizable anymore"
        }
    }
    std::vector<int32_t> col(n);
    CKI(cudaMemcpy(col.data(), cur, n * sizeof(int32_t), cudaMemcpyDeviceToHost));
    cleanup();
    if (colors) std::copy(col.begin(), col.end(), colors);
    if (categoryOffsets && categoryNodes) {
        std::vector<int32_t> cnt(nc + 1, 0);
        for (int v = 0; v < n; v++) cnt[col[v] + 1]++;
        for (int c = 0; c < nc; c++) cnt[c + 1] += cnt[c];
        std::copy(cnt.begin(), cnt.end(), categoryOffsets);
        std::vector<int32_t> at(cnt.begin(), cnt.end() - 1);
        for (int v = 0; v < n; v++) categoryNodes[at[col[v]]++] = v;
    }
    return nc;
}


// ------------------------------------------------------------------------------------------------ allocation-free variant
// gaia_b200_color_graph_device_noalloc: the speculative greedy above for callers that may not allocate (CUDA graph capture
// around them, serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// Its scratch - the two committed-colour arrays the rounds ping-pong between, the tentative colours, the
// round counters and the colour histogram - lives in ONE file-scope device array of fixed size, which every GPU that loads
// this library holds until the process exits: 3 * GAIA_NOALLOC_COLOR_MAX_NODES * 4 B + 512 B = 12 MB + 512 B.
// All calls share that array (it keeps the last call's contents; what must start clean is cleared per call), so two calls
// that overlap on one GPU race: serialise them.
namespace {

constexpr int64_t kNoAllocMaxNodes = GAIA_NOALLOC_COLOR_MAX_NODES;
constexpr int64_t kNoAllocCounterWords = 128;      // [0] nodes left, [1] overflow, [2] bad input, [64..127] histogram
constexpr int64_t kNoAllocWords = 3 * kNoAllocMaxNodes + kNoAllocCounterWords;
static __device__ __align__(256) int32_t g_noalloc_scratch[kNoAllocWords];

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noallocMutex;
int32_t* g_noallocPtr[kNoAllocMaxDevices] = {};      // device address of g_noalloc_scratch on each device (resolved once)

// host code cannot use the array's name as an address: ask the runtime, once per device, with that device current
int noalloc_scratch(int device, int32_t** out) {
    std::lock_guard<std::mutex> lock(g_noallocMutex);
    if (!g_noallocPtr[device]) {
        void* p = nullptr;
        if (cudaGetSymbolAddress(&p, g_noalloc_scratch) != cudaSuccess || !p) { cudaGetLastError(); return GAIA_ERR_CUDA; }
        g_noallocPtr[device] = (int32_t*)p;
    }
    *out = g_noallocPtr[device];
    return GAIA_OK;
}

// the CSR arrays come from the caller's device memory: offsets must start at 0 and ascend, neighbours must be other nodes
__global__ void k_check_csr(int n, const int32_t* __restrict__ off, const int32_t* __restrict__ adj, int32_t* bad) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const int b = off[v], e = off[v + 1];
    if ((v == 0 && b != 0) || b < 0 || e < b) { *bad = 1; return; }
    for (int q = b; q < e; q++) { const int u = adj[q]; if (u < 0 || u >= n || u == v) { *bad = 1; return; } }
}

void noalloc_message(char* msg, int32_t cap, const char* text) {
    if (msg && cap > 0) std::snprintf(msg, (size_t)cap, "%s", text);
}

}  // namespace

extern "C" int gaia_b200_color_graph_device_noalloc(int device, int32_t numNodes, const int32_t* dOffsets, const int32_t* dAdjacency,
                                                    int32_t* dColors, char* errMsg, int32_t errMsgCap) {
    noalloc_message(errMsg, errMsgCap, "");
    if (numNodes < 0 || (numNodes > 0 && (!dOffsets || !dAdjacency || !dColors))) {
        noalloc_message(errMsg, errMsgCap, "color_graph_device_noalloc: null pointer or negative node count");
        return GAIA_ERR_INVALID_ARG;
    }
    if ((int64_t)numNodes > kNoAllocMaxNodes) {
        char text[200];
        std::snprintf(text, sizeof(text), "color_graph_device_noalloc: %d nodes exceed the fixed scratch of GAIA_NOALLOC_COLOR_MAX_NODES = %lld "
                      "(use gaia_b200_color_graph_device, which allocates)", (int)numNodes, (long long)kNoAllocMaxNodes);
        noalloc_message(errMsg, errMsgCap, text);
        return GAIA_ERR_OVERFLOW;
    }
    if (numNodes == 0) return 0;
    int nDev = 0;
    if (cudaGetDeviceCount(&nDev) != cudaSuccess || device < 0 || device >= nDev || device >= kNoAllocMaxDevices) {
        noalloc_message(errMsg, errMsgCap, "color_graph_device_noalloc: no such CUDA device");
        return GAIA_ERR_NO_DEVICE;
    }
    if (cudaSetDevice(device) != cudaSuccess) { noalloc_message(errMsg, errMsgCap, "color_graph_device_noalloc: cudaSetDevice failed"); return GAIA_ERR_CUDA; }
    int32_t* scratch = nullptr;
    if (noalloc_scratch(device, &scratch) != GAIA_OK) {
        noalloc_message(errMsg, errMsgCap, "color_graph_device_noalloc: cudaGetSymbolAddress of the scratch array failed");
        return GAIA_ERR_CUDA;
    }
#define CKN(x) do { if ((x) != cudaSuccess) { cudaGetLastError(); noalloc_message(errMsg, errMsgCap, "color_graph_device_noalloc: CUDA error in " #x); return GAIA_ERR_CUDA; } } while (0)
    const int n = numNodes;
    int32_t* dA = scratch;
    int32_t* dB = scratch + kNoAllocMaxNodes;
    int32_t* dT = scratch + 2 * kNoAllocMaxNodes;
    int32_t* dCnt = scratch + 3 * kNoAllocMaxNodes;        // [0] left, [1] overflow, [2] bad input
    int32_t* dSizes = dCnt + 64;
    const int B = 256, G = (n + B - 1) / B;
    // the array still holds the previous call's colours: nothing is committed yet, all counters zero
    CKN(cudaMemsetAsync(dA, 0xff, (size_t)n * sizeof(int32_t), 0));
    CKN(cudaMemsetAsync(dCnt, 0, kNoAllocCounterWords * sizeof(int32_t), 0));
    k_check_csr<<<G, B>>>(n, dOffsets, dAdjacency, dCnt + 2);
    int32_t h[3] = {0, 0, 0};
    CKN(cudaMemcpy(h, dCnt, 3 * sizeof(int32_t), cudaMemcpyDeviceToHost));
    if (h[2]) { noalloc_message(errMsg, errMsgCap, "color_graph_device_noalloc: offsets do not ascend from 0, or a neighbour is out of range / the node itself"); return GAIA_ERR_INVALID_ARG; }
    int32_t *cur = dA, *nxt = dB;
    for (int round = 0; round < 10000; round++) {
        CKN(cudaMemsetAsync(dCnt, 0, 2 * sizeof(int32_t), 0));
        k_speculate<<<G, B>>>(n, dOffsets, dAdjacency, cur, dT, dCnt + 1);
        k_resolve<<<G, B>>>(n, dOffsets, dAdjacency, cur, dT, nxt, dCnt);
        CKN(cudaMemcpy(h, dCnt, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost));
        std::swap(cur, nxt);
        if (h[1]) { noalloc_message(errMsg, errMsgCap, "color_graph_device_noalloc: more than 64 colours needed"); return GAIA_ERR_UNSUPPORTED; }
        if (!h[0]) break;
    }
    if (h[0]) { noalloc_message(errMsg, errMsgCap, "color_graph_device_noalloc: the speculative rounds did not finish"); return GAIA_ERR_CUDA; }
    k_histogram<<<G, B>>>(n, cur, dSizes);
    CKN(cudaMemcpyAsync(dColors, cur, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToDevice, 0));
    int32_t sizes[64];
    CKN(cudaMemcpy(sizes, dSizes, 64 * sizeof(int32_t), cudaMemcpyDeviceToHost));
    CKN(cudaGetLastError());
#undef CKN
    int nc = 0;
    for (int c = 0; c < 64; c++) if (sizes[c]) nc = c + 1;
    return nc;
}
