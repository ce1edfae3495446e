// oracle/ref_shim.cpp -- TEST INFRASTRUCTURE ONLY.
// extern "C" door onto the reference's own Boruvka (compiled from where it lies under /root/reference, never copied):
//   /root/reference/code/utils/TreeEnergyLoss/kernels/lib_tree_filter/src/mst/boruvka.{cpp,hpp}
// The batch loop below is what mst.cu:86-117 does around it: fill a Graph per image, one std::thread per image, join.
#include <thread>
#include <vector>
#include "boruvka.hpp"

static void one_image(const int *edge_index, const float *edge_weight, int *edge_out, int V, int E) {
    Graph *g = createGraph(V, E);
    for (int i = 0; i < E; ++i) {
        g->edge[i].src = edge_index[2 * i];
        g->edge[i].dest = edge_index[2 * i + 1];
        g->edge[i].weight = edge_weight[i];
    }
    boruvkaMST(g, edge_out);
    delete[] g->edge;
    delete g;
}

extern "C" int ref_mst_forward(int B, int E, int V, const int *edge_index, const float *edge_weight, int *edge_out) {
    std::vector<std::thread> th;
    th.reserve(B);
    for (int b = 0; b < B; ++b)
        th.emplace_back(one_image, edge_index + (size_t)b * E * 2, edge_weight + (size_t)b * E,
                        edge_out + (size_t)b * (V - 1) * 2, V, E);
    for (auto &t : th) t.join();
    return 0;
}
