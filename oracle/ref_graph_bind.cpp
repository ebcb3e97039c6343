// TEST INFRASTRUCTURE ONLY -- pybind11 front end that exposes three functions of the REFERENCE's own
// C++ (/root/reference/csrc/cpu/graph_proc.cpp, compiled from where it lies; nothing is copied)
// so that oracle/make_golden_graph.py can pin our versions of them:
//   erode_mesh   graph_proc.cpp:17-77,  sample_nodes  graph_proc.cpp:79-136, and the three below.
// The reference binds all 26 functions in csrc/main.cpp:7-40; only the ones on the "next" rows of
// SURVEY.md section 8 (f2) are needed here.  Built by `make -C oracle ref` into oracle/_ref/.
#include "cpu/graph_proc.h"

PYBIND11_MODULE(ref_graph, m) {
    m.def("erode_mesh", &graph_proc::erode_mesh, "");
    m.def("sample_nodes", &graph_proc::sample_nodes, "");
    m.def("compute_edges_geodesic", &graph_proc::compute_edges_geodesic, "");   // graph_proc.cpp:155-300
    m.def("node_and_edge_clean_up", &graph_proc::node_and_edge_clean_up, "");   // graph_proc.cpp:388-438
    m.def("compute_clusters", &graph_proc::compute_clusters, "");               // graph_proc.cpp:440-481
    m.def("compute_pixel_anchors_euclidean", &graph_proc::compute_pixel_anchors_euclidean, "");   // graph_proc.cpp:610-709
}
