// TEST INFRASTRUCTURE: the reference's SimpleAligner.cpp needs blasr_libcpp, which is a
// missing submodule (/root/reference/.SUBMODULES.json).  This stub satisfies the linker;
// `-a` aborts in the oracle build.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <iostream>
#include "Alignment.hpp"
#include "SimpleAligner.hpp"
SimpleAligner::SimpleAligner() {}
void SimpleAligner::align(dagcon::Alignment&) {
    std::fprintf(stderr, "oracle build: -a (SimpleAligner/blasr_libcpp) unavailable\n");
    std::abort();
}
void SimpleAligner::operator()(dagcon::Alignment& a) { align(a); }
