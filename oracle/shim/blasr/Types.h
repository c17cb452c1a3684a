// TEST INFRASTRUCTURE: stub of the blasr_libcpp declarations SimpleAligner.hpp names.
// blasr_libcpp (pinned 9b4faa3efa2d59484af3c3562a3af6ce421e0294) is a MISSING submodule of
// the reference; the `-a` path cannot be built.  These stubs only let main.cpp parse.
#pragma once
#include <cstddef>
typedef unsigned char Nucleotide;
struct TupleMetrics { int tupleSize; void Initialize(int k) { tupleSize = k; } };
struct DNASequence { Nucleotide* seq; size_t length; DNASequence() : seq(0), length(0) {} };
struct FASTQSequence : DNASequence {};
static const int SMRTDistanceMatrix[5][5] = {{0}};
template <class A, class B> struct DistanceMatrixScoreFunction {
    int del, ins;
    void InitializeScoreMatrix(const int (*)[5]) {}
};
