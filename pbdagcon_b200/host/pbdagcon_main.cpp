// pbdagcon_main.cpp -- drop-in `pbdagcon` command line over the GPU consensus path.
// Same flags, defaults and output as the reference CLI (reference src/cpp/main.cpp:178-225 parseArgs,
// :139-145 FASTA record format): -j/--threads 4, -c/--min-coverage 6, -m/--min-length 500, -t/--trim 50,
// -a/--align, -v/--verbose, one positional input ("-" = stdin).  Records are written in input-target order
// (the reference's order is nondeterministic for -j >= 2; SURVEY.md 8(c) compares canonically).
// Extra flags: --gpus N, --batch-targets N, --batch-mb N, --stats.
#include <unistd.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "dagcon_host.hpp"

namespace {

void usage(FILE* f) {
    fputs("USAGE:\n"
          "   pbdagcon  [-v] [-a] [-t <uint>] [-m <uint>] [-c <uint>] [-j <int>] [--gpus <int>] [--stats]\n"
          "             [--] [--version] [-h] <either file path or stdin>\n\n"
          "Where:\n"
          "   -v,  --verbose             Turns on verbose logging\n"
          "   -a,  --align               Align sequences before adding to consensus\n"
          "   -t <uint>,  --trim <uint>  Trim alignments on either size (default 50)\n"
          "   -m <uint>,  --min-length <uint>    Minimum length for correction (default 500)\n"
          "   -c <uint>,  --min-coverage <uint>  Minimum coverage for correction (default 6)\n"
          "   -j <int>,  --threads <int> Number of host packing threads (reference: consensus threads; default 4)\n"
          "   --gpus <int>               GPUs to use (default: all visible)\n"
          "   --batch-targets <int>, --batch-mb <int>   targets / MiB of alignment text per GPU batch\n"
          "   --stats                    print a JSON throughput line on stderr\n"
          "   --host-align               with -a: align the 'pre' records on the host threads instead of on the GPU\n"
          "   --ascii-wire               send the alignment strings to the GPU as ASCII (default: 4-bit codes, half the bytes)\n"
          "   <either file path or stdin>  (required)  Input data (BLASR -m5 sorted by target; '-' = stdin)\n\n"
          "   PBI consensus module, B200 build\n",
          f);
}

bool parse_ull(const char* s, unsigned long long& v) {
    if (!s || !*s) return false;
    char* e = nullptr;
    v = strtoull(s, &e, 10);
    return e && *e == 0 && s[0] != '-';
}

}  // namespace

int main(int argc, char** argv) {
    dgh::ProgramOpts o;
    bool have_input = false, only_positional = false;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        if (only_positional || a == "-" || a[0] != '-') {
            if (have_input) { fprintf(stderr, "PARSE ERROR: Argument: %s\n             Too many positional arguments\n", a.c_str()); return 1; }
            o.input = a; have_input = true;
            continue;
        }
        if (a == "--") { only_positional = true; continue; }
        if (a == "-h" || a == "--help") { usage(stdout); return 0; }
        if (a == "--version") { puts("\npbdagcon  version: 0.3-b200\n"); return 0; }
        if (a == "-a" || a == "--align") { o.align = true; continue; }
        if (a == "-v" || a == "--verbose") { o.verbose = true; continue; }
        if (a == "--stats") { o.stats = true; continue; }
        if (a == "--ascii-wire") { o.asciiWire = true; continue; }
        if (a == "--host-align") { o.hostAlign = true; continue; }
        // value flags: "-j 4", "-j4", "--threads 4", "--threads=4"
        std::string name = a, val;
        bool has_val = false;
        if (a.size() > 2 && a[1] != '-') { name = a.substr(0, 2); val = a.substr(2); has_val = true; }
        else if (a.compare(0, 2, "--") == 0) {
            const size_t eq = a.find('=');
            if (eq != std::string::npos) { name = a.substr(0, eq); val = a.substr(eq + 1); has_val = true; }
        }
        const bool known = name == "-j" || name == "--threads" || name == "-c" || name == "--min-coverage" || name == "-m" ||
                           name == "--min-length" || name == "-t" || name == "--trim" || name == "--gpus" ||
                           name == "--batch-targets" || name == "--batch-mb";
        if (!known) { fprintf(stderr, "PARSE ERROR: Argument: %s\n             Couldn't find match for argument\n", a.c_str()); return 1; }
        if (!has_val) {
            if (i + 1 >= argc) { fprintf(stderr, "PARSE ERROR: Argument: %s\n             Missing a value for this argument!\n", name.c_str()); return 1; }
            val = argv[++i];
        }
        unsigned long long v = 0;
        if (!parse_ull(val.c_str(), v)) { fprintf(stderr, "PARSE ERROR: Argument: %s\n             Couldn't read argument value from string '%s'\n", name.c_str(), val.c_str()); return 1; }
        if (name == "-j" || name == "--threads") o.threads = (int)v;
        else if (name == "-c" || name == "--min-coverage") o.minCov = (size_t)v;
        else if (name == "-m" || name == "--min-length") o.minLen = (size_t)v;
        else if (name == "-t" || name == "--trim") o.trim = (unsigned)v;
        else if (name == "--gpus") o.gpus = (int)v;
        else if (name == "--batch-targets") o.batchTargets = (size_t)std::max(1ull, v);
        else if (name == "--batch-mb") o.batchBytes = (size_t)std::max(1ull, v) << 20;
    }
    if (!have_input) { fputs("PARSE ERROR: \n             Required argument missing: input\n\n", stderr); usage(stderr); return 1; }

    dgh::RunStats st;
    std::string err;
    if (o.verbose) fprintf(stderr, "pbdagcon-b200: input %s, packing threads %d\n", o.input.c_str(), o.threads);
    const int rc = dgh::run_text_pipeline(o, stdout, st, err);
    if (rc) { fprintf(stderr, "pbdagcon-b200: %s\n", err.c_str()); return 1; }
    if (o.stats || o.verbose)
        fprintf(stderr,
                "{\"targets\": %llu, \"targets_dropped\": %llu, \"alignments\": %llu, \"input_bytes\": %llu, \"batches\": %llu, "
                "\"records\": %llu, \"bases\": %llu, \"seconds\": %.4f, \"scan_seconds\": %.4f, \"scan_cut_seconds\": %.4f, \"pack_thread_seconds\": %.4f, "
                "\"pin_seconds\": %.4f, \"mbp_per_s\": %.3f, \"targets_per_s\": %.1f}\n",
                (unsigned long long)st.targets_in, (unsigned long long)st.targets_dropped, (unsigned long long)st.alignments,
                (unsigned long long)st.input_bytes, (unsigned long long)st.batches, (unsigned long long)st.records,
                (unsigned long long)st.bases, st.seconds, st.scan_seconds, st.scan_cut_seconds, st.pack_seconds, st.pin_seconds,
                st.bases / 1e6 / (st.seconds > 0 ? st.seconds : 1),
                st.targets_in / (st.seconds > 0 ? st.seconds : 1));
    // Every handle is destroyed and every record written: leave without the CUDA runtime's own process teardown
    // (unmapping the page-locked ring and the context costs several hundred ms that buy nothing at exit).
    if (fflush(stdout) != 0) { fputs("pbdagcon-b200: write error on the FASTA output\n", stderr); _exit(1); }
    fflush(stderr);
    _exit(0);
}
