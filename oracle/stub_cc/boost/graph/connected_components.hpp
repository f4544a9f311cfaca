// Stand-in for <boost/graph/connected_components.hpp>: see ../../README.md.  Not Boost code.
#pragma once
#include <vector>

#include "adjacency_list.hpp"

namespace boost {
// Writes the component number of every vertex into c[v] and returns the number of components.
// Numbering as Boost's: a depth-first search is started from each not-yet-visited vertex in
// index order, and each start opens the next component number.
template <class G, class ComponentMap>
int connected_components(const G& g, ComponentMap c) {
    const std::size_t n = g.adjacency.size();
    std::vector<char> seen(n, 0);
    std::vector<std::size_t> stack;
    int count = 0;
    for (std::size_t s = 0; s < n; ++s) {
        if (seen[s]) continue;
        seen[s] = 1;
        stack.push_back(s);
        while (!stack.empty()) {
            const std::size_t v = stack.back();
            stack.pop_back();
            c[v] = count;
            for (std::size_t w : g.adjacency[v])
                if (!seen[w]) {
                    seen[w] = 1;
                    stack.push_back(w);
                }
        }
        ++count;
    }
    return count;
}
}  // namespace boost
