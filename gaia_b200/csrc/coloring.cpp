// Vertex graph colouring for tet and triangle meshes — host code, no device work.
//
// Restates the reference's offline colouring tool (S/GraphColoring/GraphColoring.cpp with `-v -a greedy [-b -r ratio]`):
//   OrderedGreedy::color / nextNode / reduceDegree        M/GraphColoring/gready.cpp:16-117
//   GraphColor::convertToColoredCategories                M/GraphColoring/ColoringAlgorithms.cpp:69-80
//   GraphColor::balanceColoredCategories and its helpers  M/GraphColoring/ColoringAlgorithms.cpp:82-198
//   TetMeshVertexGraph / TriMeshVertexGraph::fromMesh     M/GraphColoring/TetMeshVertexGraph.h, TriMeshVertexGraph.h
// Same node order (smallest remaining degree first, lowest id on ties), same colour choice (smallest colour no coloured
// neighbour holds), same balancing moves in the same order (parity with the tool's `*.vertexColoring.json` is unpinned
// this round: tests/test_coloring.py checks validity and balance only) — but the node selection is a bucket queue instead of a scan over all nodes per step (the tool is
// O(V^2): minutes on a 73 k-node graph; this is O(E log V)).
#include <algorithm>
#include <cstdint>
#include <functional>
#include <queue>
#include <vector>

#include "../../include/gaia_b200.h"

namespace {

using Adj = std::vector<std::vector<int32_t>>;

// OrderedGreedy: returns the number of colours
int greedy_color(const Adj& graph, std::vector<int32_t>& colors) {
    const int32_t n = (int32_t)graph.size();
    colors.assign(n, -1);
    std::vector<int32_t> deg(n);
    int32_t maxDeg = 0;
    for (int32_t i = 0; i < n; i++) { deg[i] = (int32_t)graph[i].size(); maxDeg = std::max(maxDeg, deg[i]); }
    // bucket[d]: min-heap of node ids whose remaining degree was d when pushed; entries go stale when the degree drops
    using Heap = std::priority_queue<int32_t, std::vector<int32_t>, std::greater<int32_t>>;
    std::vector<Heap> bucket(maxDeg + 1);
    for (int32_t i = 0; i < n; i++) bucket[deg[i]].push(i);
    int32_t cur = 0, maxColor = -1;
    std::vector<uint8_t> used;
    for (int32_t done = 0; done < n; done++) {
        int32_t node = -1;
        for (;;) {                                  // nextNode(): smallest remaining degree, lowest id
            while (bucket[cur].empty()) cur++;
            const int32_t c = bucket[cur].top();
            bucket[cur].pop();
            if (deg[c] == cur) { node = c; break; }
        }
        used.assign((size_t)maxColor + 2, 0);
        for (int32_t nb : graph[node]) if (colors[nb] >= 0) used[colors[nb]] = 1;
        int32_t pick = 0;
        while (pick <= maxColor && used[pick]) pick++;
        if (pick > maxColor) maxColor = pick;
        colors[node] = pick;
        deg[node] = -1;                             // reduceDegree()
        for (int32_t nb : graph[node])
            if (deg[nb] != -1) {
                deg[nb]--;
                bucket[deg[nb]].push(nb);
                if (deg[nb] < cur) cur = deg[nb];
            }
    }
    return maxColor + 1;
}

// Mcs::color (M/GraphColoring/mcs.cpp:14-168): maximum cardinality search order - repeatedly the remaining node with the
// most already-ordered neighbours, lowest id on ties (the tool scans its node list front to back and keeps the first
// maximum) - then every node, in that order, takes the smallest colour none of its coloured neighbours holds. The tool
// rescans all remaining nodes per step (O(V^2)); here bucket[w] is a min-heap of the node ids that reached weight w.
int mcs_color(const Adj& graph, std::vector<int32_t>& colors) {
    const int32_t n = (int32_t)graph.size();
    colors.assign(n, -1);
    std::vector<int32_t> weight(n, 0), order;
    std::vector<uint8_t> taken(n, 0);
    order.reserve(n);
    using Heap = std::priority_queue<int32_t, std::vector<int32_t>, std::greater<int32_t>>;
    std::vector<Heap> bucket(1);
    for (int32_t i = 0; i < n; i++) bucket[0].push(i);
    int32_t cur = 0;
    for (int32_t done = 0; done < n; done++) {
        int32_t node = -1;
        for (;;) {
            while (bucket[cur].empty()) cur--;
            const int32_t c = bucket[cur].top();
            bucket[cur].pop();
            if (!taken[c] && weight[c] == cur) { node = c; break; }
        }
        taken[node] = 1;
        order.push_back(node);
        for (int32_t nb : graph[node]) {
            // (the tool also raises the weight of nodes that are already ordered; they are never looked at again)
            weight[nb]++;
            if (taken[nb]) continue;
            if ((size_t)weight[nb] >= bucket.size()) bucket.resize((size_t)weight[nb] + 1);
            bucket[weight[nb]].push(nb);
            if (weight[nb] > cur) cur = weight[nb];
        }
    }
    int32_t maxColor = -1;
    std::vector<uint8_t> used;
    for (int32_t node : order) {
        used.assign((size_t)maxColor + 2, 0);
        for (int32_t nb : graph[node]) if (colors[nb] >= 0) used[colors[nb]] = 1;
        int32_t pick = 0;
        while (pick <= maxColor && used[pick]) pick++;
        if (pick > maxColor) maxColor = pick;
        colors[node] = pick;
    }
    return maxColor + 1;
}

void balance(const Adj& graph, std::vector<int32_t>& colors, std::vector<std::vector<int32_t>>& cat, float goalMaxMinRatio) {
    if (cat.empty()) return;
    auto changable = [&](int32_t node, int32_t dst) {
        for (int32_t nb : graph[node]) if (colors[nb] == dst) return false;
        return true;
    };
    auto findChangable = [&](int src, int dst) {
        const auto& c = cat[src];
        for (size_t i = 0; i < c.size(); i++) if (changable(c[i], dst)) return (int64_t)i;
        return (int64_t)-1;
    };
    float ratio;
    // (the tool's loop has no bound and can shuttle a node between two colours for ever, e.g. with isolated nodes; the
    //  moves are capped here - a colouring is valid after every move, so stopping early only leaves it less balanced)
    int64_t moves = 0, cap = 4 * (int64_t)graph.size() + 16;
    do {
        if (++moves > cap) return;
        int big = 0, small = 0;
        size_t maxSize = cat[0].size(), minSize = cat[0].size();
        for (size_t c = 0; c < cat.size(); c++) {
            if (maxSize < cat[c].size()) { big = (int)c; maxSize = cat[c].size(); }
            if (minSize > cat[c].size()) { small = (int)c; minSize = cat[c].size(); }
        }
        ratio = float(cat[big].size()) / float(cat[small].size());
        int64_t id = findChangable(big, small);
        if (id == -1)
            for (size_t c = 0; c < cat.size(); c++) {
                if ((int)c == big || (int)c == small) continue;
                id = findChangable((int)c, small);
                if (id != -1) { big = (int)c; break; }
            }
        if (id == -1) return;                       // "not optimizable anymore"
        const int32_t node = cat[big][(size_t)id];
        colors[node] = small;
        cat[big].erase(cat[big].begin() + id);
        cat[small].push_back(node);
    } while (ratio > goalMaxMinRatio);
}

// flags: GAIA_COLOR_BALANCE | GAIA_COLOR_MCS
int color_adj(const Adj& graph, int flags, float ratio, int32_t* colors, int32_t* catOff, int32_t* catNodes) {
    const int balanceOn = flags & GAIA_COLOR_BALANCE;
    std::vector<int32_t> col;
    int nc = (flags & GAIA_COLOR_MCS) ? mcs_color(graph, col) : greedy_color(graph, col);
    std::vector<std::vector<int32_t>> cat(nc);
    for (int32_t v = 0; v < (int32_t)graph.size(); v++) cat[col[v]].push_back(v);
    if (balanceOn) balance(graph, col, cat, ratio);
    if (colors) std::copy(col.begin(), col.end(), colors);
    if (catOff && catNodes) {
        int32_t run = 0;
        for (int c = 0; c < nc; c++) {
            catOff[c] = run;
            std::copy(cat[c].begin(), cat[c].end(), catNodes + run);
            run += (int32_t)cat[c].size();
        }
        catOff[nc] = run;
    }
    return nc;
}

// unique undirected edges -> adjacency in edge order (GraphColor::GraphColor(const Graph&), ColoringAlgorithms.cpp:11-25)
void adj_from_pairs(int32_t n, std::vector<std::pair<int32_t, int32_t>>& e, Adj& graph) {
    for (auto& p : e) if (p.first > p.second) std::swap(p.first, p.second);
    std::sort(e.begin(), e.end());
    e.erase(std::unique(e.begin(), e.end()), e.end());
    graph.assign(n, {});
    for (auto& p : e) { graph[p.first].push_back(p.second); graph[p.second].push_back(p.first); }
}

}  // namespace

extern "C" {

int gaia_b200_color_graph(int32_t numNodes, int64_t nEdges, const int32_t* edges, int32_t balanceOn, float goalMaxMinRatio,
                          int32_t* colors, int32_t* categoryOffsets, int32_t* categoryNodes) {
    if (numNodes < 0 || nEdges < 0 || (nEdges > 0 && !edges)) return GAIA_ERR_INVALID_ARG;
    if (numNodes == 0) return 0;
    Adj graph(numNodes);
    for (int64_t i = 0; i < nEdges; i++) {
        const int32_t a = edges[2 * i], b = edges[2 * i + 1];
        if (a < 0 || b < 0 || a >= numNodes || b >= numNodes || a == b) return GAIA_ERR_INVALID_ARG;
        graph[a].push_back(b); graph[b].push_back(a);
    }
    return color_adj(graph, balanceOn, goalMaxMinRatio, colors, categoryOffsets, categoryNodes);
}

int gaia_b200_color_tet_mesh(int32_t nV, int32_t nT, const int32_t* tets, int32_t balanceOn, float goalMaxMinRatio,
                             int32_t* colors, int32_t* categoryOffsets, int32_t* categoryNodes) {
    if (nV <= 0 || nT < 0 || (nT > 0 && !tets)) return GAIA_ERR_INVALID_ARG;
    std::vector<std::pair<int32_t, int32_t>> e;
    e.reserve((size_t)nT * 6);
    for (int64_t t = 0; t < nT; t++)
        for (int a = 0; a < 4; a++)
            for (int b = a + 1; b < 4; b++) {
                const int32_t x = tets[4 * t + a], y = tets[4 * t + b];
                if (x < 0 || y < 0 || x >= nV || y >= nV || x == y) return GAIA_ERR_INVALID_ARG;
                e.emplace_back(x, y);
            }
    Adj graph;
    adj_from_pairs(nV, e, graph);
    return color_adj(graph, balanceOn, goalMaxMinRatio, colors, categoryOffsets, categoryNodes);
}

int gaia_b200_color_tri_mesh(int32_t nV, int32_t nF, const int32_t* faces, int32_t edgeEnergy, int32_t balanceOn, float goalMaxMinRatio,
                             int32_t* colors, int32_t* categoryOffsets, int32_t* categoryNodes) {
    if (nV <= 0 || nF < 0 || (nF > 0 && !faces)) return GAIA_ERR_INVALID_ARG;
    std::vector<std::pair<int32_t, int32_t>> e;
    e.reserve((size_t)nF * 4);
    // directed edge (a,b) -> opposite vertex, to pair the two wings of an interior edge (bending couples them)
    std::vector<std::pair<std::pair<int32_t, int32_t>, int32_t>> half;
    for (int64_t f = 0; f < nF; f++)
        for (int k = 0; k < 3; k++) {
            const int32_t a = faces[3 * f + k], b = faces[3 * f + (k + 1) % 3], c = faces[3 * f + (k + 2) % 3];
            if (a < 0 || b < 0 || a >= nV || b >= nV || a == b) return GAIA_ERR_INVALID_ARG;
            e.emplace_back(a, b);
            if (edgeEnergy) half.push_back({{std::min(a, b), std::max(a, b)}, c});
        }
    if (edgeEnergy) {
        std::sort(half.begin(), half.end());
        for (size_t i = 0; i + 1 < half.size(); i++)
            if (half[i].first == half[i + 1].first && half[i].second != half[i + 1].second) e.emplace_back(half[i].second, half[i + 1].second);
    }
    Adj graph;
    adj_from_pairs(nV, e, graph);
    return color_adj(graph, balanceOn, goalMaxMinRatio, colors, categoryOffsets, categoryNodes);
}

}  // extern "C"
