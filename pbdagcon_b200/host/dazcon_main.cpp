// dazcon_main.cpp -- drop-in `dazcon` command line over the GPU consensus path.
// Same flags and defaults as the reference CLI (reference src/cpp/dazcon.cpp:122-192 parseArgs): -s/--seq-file (required),
// -a/--align-file (required), -j/--threads 4, -c/--min-coverage 6, -l/--min-len 500, -t/--trim 10, -m/--max-hit 85,
// -x/--coverage-sort, -o/--only-proper-overlaps, -v/--verbose, positional target ids.  Output: ">id/n/b_e\nSEQ\n" on
// stdout (:92-99), n = running record index in input order (what `dazcon -j 1` prints; the reference's counter is an
// uninitialised per-thread variable, :62).  Extra flags: --gpus N, --batch-targets N, --batch-mb N, --stats.
#include <unistd.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "dagcon_host.hpp"

namespace {

void usage(FILE* f) {
    fputs("USAGE:\n"
          "   dazcon  [-v] [-o] [-x] [-m <uint>] -s <string> -a <string> [-t <uint>] [-l <uint>] [-c <uint>] [-j <int>]\n"
          "           [--gpus <int>] [--stats] [--] [--version] [-h] <list of ints> ...\n\n"
          "Where:\n"
          "   -v,  --verbose                  Turns on verbose logging\n"
          "   -o,  --only-proper-overlaps     Use only 'proper overlaps', i.e., align to the ends\n"
          "   -x,  --coverage-sort            Sort hits by coverage\n"
          "   -m <uint>,  --max-hit <uint>    Maximum number of hits to pass to consensus (default 85)\n"
          "   -s <string>,  --seq-file <string>    (required)  Path to the sequences file\n"
          "   -a <string>,  --align-file <string>  (required)  Path to the alignments file\n"
          "   -t <uint>,  --trim <uint>       Trim alignments on either size (default 10)\n"
          "   -l <uint>,  --min-len <uint>    Minimum length for correction (default 500)\n"
          "   -c <uint>,  --min-coverage <uint>  Minimum coverage for correction (default 6)\n"
          "   -j <int>,  --threads <int>      Number of host realignment threads (reference: consensus threads; default 4)\n"
          "   --gpus <int>                    GPUs to use (default: all visible)\n"
          "   --batch-targets <int>, --batch-mb <int>   targets / MiB of alignment columns per GPU batch\n"
          "   --stats                         print a JSON throughput line on stderr\n"
          "   --host-realign                  rebuild the alignments from the trace points on the host threads (dazzler's\n"
          "                                   Compute_Trace_PTS) instead of on the GPU\n"
          "   <list of ints>                  Limit consensus to list of target ids\n\n"
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
    o.trim = 10;            // dazcon's default (dazcon.cpp:143-146); pbdagcon's is 50
    o.batchTargets = 4096;
    o.batchBytes = 2048ull << 20;
    bool only_positional = false;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        if (only_positional || a[0] != '-' || a == "-") {
            unsigned long long v = 0;
            if (!parse_ull(a.c_str(), v)) { fprintf(stderr, "PARSE ERROR: Argument: (--targets)\n             Couldn't read argument value from string '%s'\n", a.c_str()); return 1; }
            o.targets.insert((int)v);
            continue;
        }
        if (a == "--") { only_positional = true; continue; }
        if (a == "-h" || a == "--help") { usage(stdout); return 0; }
        if (a == "--version") { puts("\ndazcon  version: 0.3-b200\n"); return 0; }
        if (a == "-v" || a == "--verbose") { o.verbose = true; continue; }
        if (a == "-x" || a == "--coverage-sort") { o.sortCov = true; continue; }
        if (a == "-o" || a == "--only-proper-overlaps") { o.properOvls = true; continue; }
        if (a == "--stats") { o.stats = true; continue; }
        if (a == "--host-realign") { o.hostRealign = true; continue; }
        // combined switches as TCLAP accepts them: -ox, -xo, -vox ...
        if (a.size() > 2 && a[1] != '-' && a.find_first_not_of("vxo", 1) == std::string::npos) {
            for (size_t k = 1; k < a.size(); k++) {
                if (a[k] == 'v') o.verbose = true;
                if (a[k] == 'x') o.sortCov = true;
                if (a[k] == 'o') o.properOvls = true;
            }
            continue;
        }
        std::string name = a, val;
        bool has_val = false;
        if (a.size() > 2 && a[1] != '-') { name = a.substr(0, 2); val = a.substr(2); has_val = true; }
        else if (a.compare(0, 2, "--") == 0) {
            const size_t eq = a.find('=');
            if (eq != std::string::npos) { name = a.substr(0, eq); val = a.substr(eq + 1); has_val = true; }
        }
        const bool str_flag = name == "-s" || name == "--seq-file" || name == "-a" || name == "--align-file";
        const bool known = str_flag || name == "-j" || name == "--threads" || name == "-c" || name == "--min-coverage" || name == "-l" ||
                           name == "--min-len" || name == "-t" || name == "--trim" || name == "-m" || name == "--max-hit" ||
                           name == "--gpus" || name == "--batch-targets" || name == "--batch-mb";
        if (!known) { fprintf(stderr, "PARSE ERROR: Argument: %s\n             Couldn't find match for argument\n", a.c_str()); return 1; }
        if (!has_val) {
            if (i + 1 >= argc) { fprintf(stderr, "PARSE ERROR: Argument: %s\n             Missing a value for this argument!\n", name.c_str()); return 1; }
            val = argv[++i];
        }
        if (name == "-s" || name == "--seq-file") { o.seqFile = val; continue; }
        if (name == "-a" || name == "--align-file") { o.alnFile = val; continue; }
        unsigned long long v = 0;
        if (!parse_ull(val.c_str(), v)) { fprintf(stderr, "PARSE ERROR: Argument: %s\n             Couldn't read argument value from string '%s'\n", name.c_str(), val.c_str()); return 1; }
        if (name == "-j" || name == "--threads") o.threads = (int)v;
        else if (name == "-c" || name == "--min-coverage") o.minCov = (size_t)v;
        else if (name == "-l" || name == "--min-len") o.minLen = (size_t)v;
        else if (name == "-t" || name == "--trim") o.trim = (unsigned)v;
        else if (name == "-m" || name == "--max-hit") o.maxHits = (unsigned)v;
        else if (name == "--gpus") o.gpus = (int)v;
        else if (name == "--batch-targets") o.batchTargets = (size_t)std::max(1ull, v);
        else if (name == "--batch-mb") o.batchBytes = (size_t)std::max(1ull, v) << 20;
    }
    if (o.alnFile.empty()) { fputs("PARSE ERROR: \n             Required argument missing: align-file\n\n", stderr); usage(stderr); return 1; }
    if (o.seqFile.empty()) { fputs("PARSE ERROR: \n             Required argument missing: seq-file\n\n", stderr); usage(stderr); return 1; }

    dgh::RunStats st;
    std::string err;
    if (o.verbose) fprintf(stderr, "dazcon-b200: db %s, las %s, realignment threads %d\n", o.seqFile.c_str(), o.alnFile.c_str(), o.threads);
    const int rc = dgh::run_daz_pipeline(o, stdout, st, err);
    if (rc) { fprintf(stderr, "dazcon-b200: %s\n", err.c_str()); return 1; }
    if (o.stats || o.verbose)
        fprintf(stderr,
                "{\"targets\": %llu, \"targets_dropped\": %llu, \"alignments\": %llu, \"batches\": %llu, "
                "\"records\": %llu, \"bases\": %llu, \"seconds\": %.4f, \"scan_seconds\": %.4f, \"mbp_per_s\": %.3f, \"targets_per_s\": %.1f}\n",
                (unsigned long long)st.targets_in, (unsigned long long)st.targets_dropped, (unsigned long long)st.alignments,
                (unsigned long long)st.batches, (unsigned long long)st.records,
                (unsigned long long)st.bases, st.seconds, st.scan_seconds, st.bases / 1e6 / (st.seconds > 0 ? st.seconds : 1),
                st.targets_in / (st.seconds > 0 ? st.seconds : 1));
    if (fflush(stdout) != 0) { fputs("dazcon-b200: write error on the FASTA output\n", stderr); _exit(1); }
    fflush(stderr);
    _exit(0);
}
