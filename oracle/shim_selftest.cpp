// TEST INFRASTRUCTURE (oracle/): self-test of the Boost.Graph stand-in (oracle/shim/boost/graph/adjacency_list.hpp).
// Every tie-break of the reference rests on the ordering rules of adjacency_list<vecS, vecS, bidirectionalS>
// (SURVEY.md 8(c) "residual assumption", App. B.1); this program pins the shim to those documented rules:
//   add_vertex = next dense id; add_edge appends to out(u) and in(v) and never rejects a parallel edge;
//   edge(u,v) = FIRST match in out(u); clear_vertex erases order-preservingly from the neighbours' lists;
//   edge properties keep their address (and value) while other edges are removed.
// Prints "shim ok" and exits 0, or names the failed rule.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <boost/graph/adjacency_list.hpp>

struct VP { int tag = 0; };
struct EP { int count = 0; bool visited = false; };
typedef boost::adjacency_list<boost::vecS, boost::vecS, boost::bidirectionalS, VP, EP> G;
typedef boost::graph_traits<G>::edge_descriptor ED;

#define CHECK(c) do { if (!(c)) { printf("shim FAILED: %s (line %d)\n", #c, __LINE__); exit(1); } } while (0)

static std::vector<size_t> outs(size_t v, G& g) {
    std::vector<size_t> r;
    boost::graph_traits<G>::out_edge_iterator a, b;
    for (boost::tie(a, b) = boost::out_edges(v, g); a != b; ++a) r.push_back(boost::target(*a, g));
    return r;
}
static std::vector<size_t> ins(size_t v, G& g) {
    std::vector<size_t> r;
    boost::graph_traits<G>::in_edge_iterator a, b;
    for (boost::tie(a, b) = boost::in_edges(v, g); a != b; ++a) r.push_back(boost::source(*a, g));
    return r;
}

int main() {
    G g(4);
    CHECK(boost::num_vertices(g) == 4);
    CHECK(boost::add_vertex(g) == 4 && boost::add_vertex(g) == 5);          // dense ids, next index
    // append order in out(u) and in(v)
    boost::add_edge(0, 3, g); boost::add_edge(0, 1, g); boost::add_edge(0, 2, g);
    boost::add_edge(1, 3, g); boost::add_edge(2, 3, g); boost::add_edge(4, 3, g);
    CHECK((outs(0, g) == std::vector<size_t>{3, 1, 2}));
    CHECK((ins(3, g) == std::vector<size_t>{0, 1, 2, 4}));
    // parallel edges are accepted; edge(u,v) returns the FIRST one
    ED e1 = boost::add_edge(0, 1, g).first;
    g[e1].count = 7;
    CHECK((outs(0, g) == std::vector<size_t>{3, 1, 2, 1}));
    CHECK(boost::out_degree(0, g) == 4 && boost::in_degree(1, g) == 2);
    std::pair<ED, bool> f = boost::edge(0, 1, g);
    CHECK(f.second && g[f.first].count == 0);                                // the first (0,1), not the count-7 one
    g[f.first].count = 3; g[f.first].visited = true;
    CHECK(!boost::edge(1, 0, g).second && !boost::edge(5, 0, g).second);
    // properties have stable addresses across removals of other edges
    EP* keep = &g[f.first];
    ED e23 = boost::edge(2, 3, g).first;
    g[e23].count = 11;
    // clear_vertex(1): removes 0->1 (both), 1->3; neighbours' lists keep the order of what remains
    boost::clear_vertex(1, g);
    CHECK(boost::out_degree(1, g) == 0 && boost::in_degree(1, g) == 0);
    CHECK((outs(0, g) == std::vector<size_t>{3, 2}));
    CHECK((ins(3, g) == std::vector<size_t>{0, 2, 4}));
    CHECK(g[boost::edge(2, 3, g).first].count == 11);
    (void)keep;
    // vertex ids survive clear_vertex (no renumbering): 4 and 5 are still there, 5 is isolated
    CHECK(boost::num_vertices(g) == 6 && boost::out_degree(4, g) == 1 && boost::out_degree(5, g) == 0);
    // clearing a vertex in the middle of an in-list
    boost::clear_vertex(2, g);
    CHECK((ins(3, g) == std::vector<size_t>{0, 4}) && (outs(0, g) == std::vector<size_t>{3}));
    // edges() enumerates what is left, in creation order
    boost::graph_traits<G>::edge_iterator a, b;
    std::vector<std::pair<size_t, size_t>> all;
    for (boost::tie(a, b) = boost::edges(g); a != b; ++a) all.push_back({boost::source(*a, g), boost::target(*a, g)});
    CHECK((all == std::vector<std::pair<size_t, size_t>>{{0, 3}, {4, 3}}));
    // add after clear: appended behind the survivors
    boost::add_edge(5, 3, g); boost::add_edge(0, 5, g);
    CHECK((ins(3, g) == std::vector<size_t>{0, 4, 5}) && (outs(0, g) == std::vector<size_t>{3, 5}));
    // vertex_index map is the identity
    CHECK(boost::get(boost::vertex_index, g)[3] == 3);
    printf("shim ok\n");
    return 0;
}
