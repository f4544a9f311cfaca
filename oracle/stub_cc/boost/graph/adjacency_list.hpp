// Stand-in for <boost/graph/adjacency_list.hpp>: see ../../README.md.  Not Boost code.
#pragma once
#include <cstddef>
#include <utility>
#include <vector>

namespace boost {
struct vecS {};
struct undirectedS {};

template <class OutEdgeList = vecS, class VertexList = vecS, class Directed = undirectedS>
class adjacency_list {
  public:
    using vertex_descriptor = std::size_t;
    struct edge_descriptor {
        std::size_t source, target;
    };
    std::vector<std::vector<std::size_t>> adjacency;  // per vertex, in insertion order
    std::vector<edge_descriptor> edge_list;            // in insertion order
};

template <class G>
std::size_t add_vertex(G& g) {
    g.adjacency.emplace_back();
    return g.adjacency.size() - 1;
}

// vecS vertex storage: adding an edge to a vertex beyond the end grows the vertex set
template <class G>
std::pair<typename G::edge_descriptor, bool> add_edge(std::size_t u, std::size_t v, G& g) {
    const std::size_t need = (u > v ? u : v) + 1;
    if (g.adjacency.size() < need) g.adjacency.resize(need);
    g.adjacency[u].push_back(v);
    if (u != v) g.adjacency[v].push_back(u);
    g.edge_list.push_back({u, v});
    return {g.edge_list.back(), true};
}

template <class G>
std::size_t num_vertices(const G& g) {
    return g.adjacency.size();
}

template <class G>
std::pair<typename std::vector<typename G::edge_descriptor>::const_iterator,
          typename std::vector<typename G::edge_descriptor>::const_iterator>
edges(const G& g) {
    return {g.edge_list.begin(), g.edge_list.end()};
}

template <class G>
std::size_t source(const typename G::edge_descriptor& e, const G&) {
    return e.source;
}
template <class G>
std::size_t target(const typename G::edge_descriptor& e, const G&) {
    return e.target;
}

template <class It>
struct iterator_range {
    It b, e;
    It begin() const { return b; }
    It end() const { return e; }
};
template <class It>
iterator_range<It> make_iterator_range(const std::pair<It, It>& p) {
    return {p.first, p.second};
}
}  // namespace boost
