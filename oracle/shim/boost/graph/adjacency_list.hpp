// TEST INFRASTRUCTURE (oracle/): a minimal stand-in for the subset of Boost.Graph
// that the reference's AlnGraphBoost.cpp uses, so that the UNMODIFIED reference
// sources under /root/reference/src/cpp compile in this container (Boost is absent
// and cannot be installed).  Written from the documented behaviour of
// adjacency_list<vecS, vecS, bidirectionalS, VP, EP>:
//   * vertex descriptors are dense size_t; add_vertex returns the next index
//   * add_edge appends to out(u) and in(v); parallel edges are never rejected
//   * edge(u,v,g) returns the FIRST match in out(u)
//   * clear_vertex erases order-preservingly from the neighbours' lists
//   * edge properties have stable addresses (global std::list of edges)
// Never linked into the product library.
#pragma once
#include <cstddef>
#include <list>
#include <vector>
#include <utility>
#include <tuple>
#include <iostream>
#include <algorithm>

namespace boost {

struct vecS {};
struct listS {};
struct bidirectionalS {};
struct no_property {};
enum vertex_index_t { vertex_index };

using std::tie;

namespace detail {
// one non-template edge descriptor shared by every instantiation, as in BGL
// (edge_desc_impl<bidirectional_tag, size_t>)
struct edge_desc_impl {
    std::size_t m_source;
    std::size_t m_target;
    void* m_eproperty;
    edge_desc_impl() : m_source(0), m_target(0), m_eproperty(nullptr) {}
    edge_desc_impl(std::size_t s, std::size_t t, void* p) : m_source(s), m_target(t), m_eproperty(p) {}
    bool operator==(const edge_desc_impl& o) const { return m_eproperty == o.m_eproperty; }
    bool operator!=(const edge_desc_impl& o) const { return m_eproperty != o.m_eproperty; }
};

struct counting_iter {
    std::size_t v;
    counting_iter(std::size_t x = 0) : v(x) {}
    std::size_t operator*() const { return v; }
    counting_iter& operator++() { ++v; return *this; }
    counting_iter operator++(int) { counting_iter t = *this; ++v; return t; }
    bool operator==(const counting_iter& o) const { return v == o.v; }
    bool operator!=(const counting_iter& o) const { return v != o.v; }
};

struct identity_index_map {
    std::size_t operator[](std::size_t v) const { return v; }
};
}  // namespace detail
// in namespace boost itself so that ADL finds make_label_writer through get()'s return type
template <class T> struct dummy_pmap {};
template <class T> struct dummy_writer {};

template <class OutEdgeListS, class VertexListS, class DirectedS,
          class VP = no_property, class EP = no_property>
class adjacency_list {
public:
    typedef std::size_t vertex_descriptor;
    typedef detail::edge_desc_impl edge_descriptor;
    struct stored_edge {
        std::size_t src, dst;
        EP prop;
        stored_edge(std::size_t s, std::size_t d) : src(s), dst(d), prop() {}
    };
    typedef std::list<stored_edge> edge_list;
    typedef typename edge_list::iterator edge_ref;
    struct vertex_rec {
        std::vector<edge_ref> out, in;
        VP prop;
        vertex_rec() : prop() {}
    };

    template <class It>
    struct edge_iter_adaptor {
        It it;
        edge_iter_adaptor() {}
        edge_iter_adaptor(It i) : it(i) {}
        edge_descriptor operator*() const {
            return edge_descriptor((*it)->src, (*it)->dst, (void*)&((*it)->prop));
        }
        edge_iter_adaptor& operator++() { ++it; return *this; }
        edge_iter_adaptor operator++(int) { edge_iter_adaptor t = *this; ++it; return t; }
        bool operator==(const edge_iter_adaptor& o) const { return it == o.it; }
        bool operator!=(const edge_iter_adaptor& o) const { return it != o.it; }
    };
    struct all_edge_iter {
        edge_ref it;
        all_edge_iter() {}
        all_edge_iter(edge_ref i) : it(i) {}
        edge_descriptor operator*() const { return edge_descriptor(it->src, it->dst, (void*)&(it->prop)); }
        all_edge_iter& operator++() { ++it; return *this; }
        bool operator==(const all_edge_iter& o) const { return it == o.it; }
        bool operator!=(const all_edge_iter& o) const { return it != o.it; }
    };

    typedef detail::counting_iter vertex_iterator;
    typedef all_edge_iter edge_iterator;
    typedef edge_iter_adaptor<typename std::vector<edge_ref>::iterator> out_edge_iterator;
    typedef edge_iter_adaptor<typename std::vector<edge_ref>::iterator> in_edge_iterator;

    adjacency_list() {}
    explicit adjacency_list(std::size_t n) : m_vertices(n) {}
    // graphs are only copied/assigned while edge-free in the reference (`_g = G(n)`)
    adjacency_list(const adjacency_list& o) : m_vertices(o.m_vertices) { }
    adjacency_list& operator=(const adjacency_list& o) {
        m_vertices = o.m_vertices; m_edges.clear(); return *this;
    }

    VP& operator[](vertex_descriptor v) { return m_vertices[v].prop; }
    const VP& operator[](vertex_descriptor v) const { return m_vertices[v].prop; }
    EP& operator[](const edge_descriptor& e) { return *static_cast<EP*>(e.m_eproperty); }

    std::vector<vertex_rec> m_vertices;
    edge_list m_edges;
};

template <class G> struct graph_traits {
    typedef typename G::vertex_descriptor vertex_descriptor;
    typedef typename G::edge_descriptor edge_descriptor;
    typedef typename G::vertex_iterator vertex_iterator;
    typedef typename G::edge_iterator edge_iterator;
    typedef typename G::in_edge_iterator in_edge_iterator;
    typedef typename G::out_edge_iterator out_edge_iterator;
};

template <class G, class Tag> struct property_map { typedef detail::identity_index_map type; };

#define BGL_SHIM_TPL template <class O, class V, class D, class VP, class EP>
#define BGL_SHIM_G adjacency_list<O, V, D, VP, EP>

BGL_SHIM_TPL std::pair<typename BGL_SHIM_G::edge_descriptor, bool>
add_edge(std::size_t u, std::size_t v, BGL_SHIM_G& g) {
    typedef BGL_SHIM_G G;
    std::size_t need = std::max(u, v) + 1;
    if (g.m_vertices.size() < need) g.m_vertices.resize(need);
    g.m_edges.push_back(typename G::stored_edge(u, v));
    typename G::edge_ref r = --g.m_edges.end();
    g.m_vertices[u].out.push_back(r);
    g.m_vertices[v].in.push_back(r);
    return std::make_pair(typename G::edge_descriptor(u, v, (void*)&r->prop), true);
}

BGL_SHIM_TPL std::size_t add_vertex(BGL_SHIM_G& g) {
    g.m_vertices.push_back(typename BGL_SHIM_G::vertex_rec());
    return g.m_vertices.size() - 1;
}

BGL_SHIM_TPL std::pair<typename BGL_SHIM_G::vertex_iterator, typename BGL_SHIM_G::vertex_iterator>
vertices(const BGL_SHIM_G& g) {
    typedef typename BGL_SHIM_G::vertex_iterator It;
    return std::make_pair(It(0), It(g.m_vertices.size()));
}

BGL_SHIM_TPL std::pair<typename BGL_SHIM_G::edge_iterator, typename BGL_SHIM_G::edge_iterator>
edges(BGL_SHIM_G& g) {
    typedef typename BGL_SHIM_G::edge_iterator It;
    return std::make_pair(It(g.m_edges.begin()), It(g.m_edges.end()));
}

BGL_SHIM_TPL std::pair<typename BGL_SHIM_G::out_edge_iterator, typename BGL_SHIM_G::out_edge_iterator>
out_edges(std::size_t v, BGL_SHIM_G& g) {
    typedef typename BGL_SHIM_G::out_edge_iterator It;
    return std::make_pair(It(g.m_vertices[v].out.begin()), It(g.m_vertices[v].out.end()));
}

BGL_SHIM_TPL std::pair<typename BGL_SHIM_G::in_edge_iterator, typename BGL_SHIM_G::in_edge_iterator>
in_edges(std::size_t v, BGL_SHIM_G& g) {
    typedef typename BGL_SHIM_G::in_edge_iterator It;
    return std::make_pair(It(g.m_vertices[v].in.begin()), It(g.m_vertices[v].in.end()));
}

BGL_SHIM_TPL std::size_t source(const detail::edge_desc_impl& e, const BGL_SHIM_G&) { return e.m_source; }
BGL_SHIM_TPL std::size_t target(const detail::edge_desc_impl& e, const BGL_SHIM_G&) { return e.m_target; }
BGL_SHIM_TPL std::size_t out_degree(std::size_t v, const BGL_SHIM_G& g) { return g.m_vertices[v].out.size(); }
BGL_SHIM_TPL std::size_t in_degree(std::size_t v, const BGL_SHIM_G& g) { return g.m_vertices[v].in.size(); }
BGL_SHIM_TPL std::size_t num_vertices(const BGL_SHIM_G& g) { return g.m_vertices.size(); }

BGL_SHIM_TPL std::pair<typename BGL_SHIM_G::edge_descriptor, bool>
edge(std::size_t u, std::size_t v, BGL_SHIM_G& g) {
    typedef typename BGL_SHIM_G::edge_descriptor ED;
    std::vector<typename BGL_SHIM_G::edge_ref>& ol = g.m_vertices[u].out;
    for (std::size_t i = 0; i < ol.size(); i++)
        if (ol[i]->dst == v) return std::make_pair(ED(u, v, (void*)&ol[i]->prop), true);
    return std::make_pair(ED(), false);
}

BGL_SHIM_TPL void clear_vertex(std::size_t u, BGL_SHIM_G& g) {
    typedef typename BGL_SHIM_G::edge_ref R;
    typename BGL_SHIM_G::vertex_rec& vr = g.m_vertices[u];
    // out-edges first (bidirectional clear_vertex = clear_out_edges + clear_in_edges)
    for (std::size_t i = 0; i < vr.out.size(); i++) {
        R r = vr.out[i];
        std::vector<R>& til = g.m_vertices[r->dst].in;
        til.erase(std::remove(til.begin(), til.end(), r), til.end());
        if (r->dst == u) continue;  // self loop: erased from the global list below, once
        g.m_edges.erase(r);
    }
    for (std::size_t i = 0; i < vr.out.size(); i++)
        if (vr.out[i]->dst == u) g.m_edges.erase(vr.out[i]);
    vr.out.clear();
    for (std::size_t i = 0; i < vr.in.size(); i++) {
        R r = vr.in[i];
        std::vector<R>& sol = g.m_vertices[r->src].out;
        sol.erase(std::remove(sol.begin(), sol.end(), r), sol.end());
        g.m_edges.erase(r);
    }
    vr.in.clear();
}

BGL_SHIM_TPL void remove_vertex(std::size_t u, BGL_SHIM_G& g) {
    // vecS semantics: erase and renumber every descriptor above u
    g.m_vertices.erase(g.m_vertices.begin() + u);
    for (typename BGL_SHIM_G::edge_ref r = g.m_edges.begin(); r != g.m_edges.end(); ++r) {
        if (r->src > u) r->src--;
        if (r->dst > u) r->dst--;
    }
}

BGL_SHIM_TPL detail::identity_index_map get(vertex_index_t, const BGL_SHIM_G&) {
    return detail::identity_index_map();
}

// graphviz stubs: printGraph() is off the consensus path; they only need to compile
template <class T, class C, class O, class V, class D, class VP, class EP>
dummy_pmap<T> get(T C::*, BGL_SHIM_G&) { return dummy_pmap<T>(); }
template <class T> dummy_writer<T> make_label_writer(dummy_pmap<T>) { return dummy_writer<T>(); }
template <class G, class W1, class W2>
void write_graphviz(std::ostream& os, const G&, W1, W2) { os << "digraph G { /* graphviz unavailable in oracle shim */ }\n"; }

#undef BGL_SHIM_TPL
#undef BGL_SHIM_G
}  // namespace boost
