// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" shim over the REFERENCE's own graph colouring: OrderedGreedy::color (Simulator/Modules/GraphColoring/gready.cpp:47-117),
// Mcs::color (mcs.cpp:14-...) and GraphColor::balanceColoredCategories (ColoringAlgorithms.cpp:152-198), compiled from where
// they lie under /root/reference (oracle/Makefile builds a tree of symlinks beside the stub Graph.h). Output: oracle/_ref/.
// balanceColoredCategories does not terminate on every graph (it may shuttle one node between two colours for ever): the
// golden generator calls it in a child process with a time limit.
#include "GraphColoring/gready.cpp"
#include "GraphColoring/mcs.cpp"
#include "GraphColoring/ColoringAlgorithms.cpp"

extern "C" {
// algorithm: 0 OrderedGreedy, 1 MCS. edges: nEdges pairs (a < b, each once), as Graph::edges holds them.
// Returns the number of colours; colours[node] in [0, nColours).
int ref_color_graph(int numNodes, long long nEdges, const int* edges, int algorithm, int balance, float goalMaxMinRatio, int* colours) {
    GAIA::GraphColoring::Graph g;
    g.numNodes = (size_t)numNodes;
    for (long long i = 0; i < nEdges; i++) g.edges.push_back({edges[2 * i], edges[2 * i + 1]});
    std::unique_ptr<GAIA::GraphColoring::GraphColor> c;
    if (algorithm == 1) c.reset(new GAIA::GraphColoring::Mcs(g)); else c.reset(new GAIA::GraphColoring::OrderedGreedy(g));
    c->color();
    if (balance) {      // the tool's order: colour, categories, balance (Simulator/GraphColoring/GraphColoring.cpp)
        c->convertToColoredCategories();
        c->balanceColoredCategories(goalMaxMinRatio);
    }
    for (int i = 0; i < numNodes; i++) colours[i] = c->get_color(i);
    return c->get_num_colors();
}
}
