// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for Simulator/Modules/GraphColoring/Graph.h when the reference's colouring
// algorithms (gready.cpp, mcs.cpp, ColoringAlgorithms.cpp) are compiled into oracle/_ref: the real header pulls in
// MeshFrame2 (mesh loaders) only to offer Graph::fromMesh, which the algorithms never call. Same members otherwise.
#pragma once
#include <array>
#include <cstddef>
#include <fstream>
#include <memory>
#include <set>
#include <string>
#include <vector>

namespace GAIA {
namespace GraphColoring {
struct Graph {
    size_t numNodes = 0;
    typedef std::array<int, 2> Edge;
    std::vector<Edge> edges;
    std::vector<float> edgeWeights;
};
}  // namespace GraphColoring
}  // namespace GAIA
