// ntcomp-b200 -- command-line front end over the C ABI (include/ntcomp_b200.h), mirroring the
// reference's three subcommands (src/cli.rs:26-93, src/main.rs:91-214):
//   ntcomp-b200 build  -o <prefix> [-k 31] [-t threads] [-l list] [--gpu [--device N]] <fastX>...     ->  <prefix>.ntcidx
//   ntcomp-b200 encode -i <prefix> <reads.fastX[.gz]>   > encoded.dat
//   ntcomp-b200 decode -i <prefix> <encoded.dat>        > decoded.fasta
// encoded.dat is the reference's format.  The index file is this library's own flat
// format (ntc_index_save): kbo's .sbwt/.lcs serialisation cannot be restated without its source, so a
// Rust host passes its loaded index through ntc_index_from_arrays instead.
// Host orchestration only: all compute goes through libntcomp_b200.so and needs a GPU (no CPU fallback).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include "../../include/ntcomp_b200.h"

static int usage() {
    fprintf(stderr,
            "usage: ntcomp-b200 build  -o <prefix> [-k 31] [-t threads] [-l input-list] [--gpu [--device N]] <fastX>...\n"
            "       ntcomp-b200 encode -i <prefix> [-t threads] [--device N] [-o out] <reads.fastX>\n"
            "       ntcomp-b200 decode -i <prefix> [-t threads] [--device N] [-o out] <encoded.dat>\n");
    return 2;
}
static int fail(const char* what, int32_t rc) {
    fprintf(stderr, "ntcomp-b200: %s failed (status %d): %s\n", what, rc, ntc_last_error());
    return 1;
}
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv) {
    if (argc < 2) return usage();
    const std::string cmd = argv[1];
    std::string prefix, out = "-", list;
    std::vector<std::string> inputs;
    uint32_t k = 31;                    // src/cli.rs:43
    int threads = 0, device = 0;
    bool gpu_build = false;
    for (int i = 2; i < argc; i++) {
        std::string a = argv[i];
        auto val = [&]() -> const char* { if (i + 1 >= argc) { usage(); exit(2); } return argv[++i]; };
        if (a == "-o" || a == "--output-prefix" || a == "--output") { std::string v = val(); if (cmd == "build") prefix = v; else out = v; }
        else if (a == "-i" || a == "--index") prefix = val();
        else if (a == "-k") k = (uint32_t)atoi(val());
        else if (a == "-t" || a == "--threads") threads = atoi(val());
        else if (a == "-l" || a == "--input-list") list = val();
        else if (a == "--device") device = atoi(val());
        else if (a == "--gpu") gpu_build = true;
        else if (a == "-p" || a == "--prefix-precalc" || a == "-m" || a == "--mem-gb" || a == "--temp-dir") (void)val();  // kbo build knobs: accepted, unused
        else if (a == "-d" || a == "--dedup-batches" || a == "--verbose") {}
        else if (a == "-h" || a == "--help") return usage();
        else if (a.size() > 1 && a[0] == '-') { fprintf(stderr, "unknown option %s\n", a.c_str()); return usage(); }
        else inputs.push_back(a);
    }
    if (prefix.empty()) return usage();
    const std::string idx_path = prefix + ".ntcidx";
    const double t0 = now();

    if (cmd == "build") {
        if (!list.empty()) {            // one path per line, or name<TAB>path (src/main.rs:64-89)
            std::ifstream f(list);
            std::string ln;
            while (std::getline(f, ln)) {
                if (ln.empty()) continue;
                size_t tab = ln.find('\t');
                inputs.push_back(tab == std::string::npos ? ln : ln.substr(tab + 1));
            }
        }
        if (inputs.empty()) return usage();
        std::vector<const char*> paths;
        for (auto& s : inputs) paths.push_back(s.c_str());
        ntc_index* idx = nullptr;
        fprintf(stderr, "Building SBWT index from %zu files...\n", inputs.size());
        int32_t rc = gpu_build ? ntc_index_build_from_files_gpu(device, paths.data(), paths.size(), k, /*add_revcomp=*/1, threads, &idx)
                               : ntc_index_build_from_files(paths.data(), paths.size(), k, /*add_revcomp=*/1, threads, &idx);   // src/main.rs:118
        if (rc) return fail("build", rc);
        fprintf(stderr, "Serializing index (k = %u, %llu nodes) to %s ...\n", ntc_index_k(idx), (unsigned long long)ntc_index_n_sets(idx),
                idx_path.c_str());
        rc = ntc_index_save(idx, idx_path.c_str());
        ntc_index_free(idx);
        if (rc) return fail("save", rc);
        return 0;
    }
    if (cmd != "encode" && cmd != "decode") return usage();
    if (inputs.size() != 1) return usage();
    ntc_index* idx = nullptr;
    fprintf(stderr, "Loading SBWT index...\n");
    int32_t rc = ntc_index_load(idx_path.c_str(), &idx);
    if (rc) return fail("index load", rc);
    if (getenv("NTC_TRACE")) fprintf(stderr, "[ntc trace] index file loaded in %.3f s\n", now() - t0);
    ntc_ctx* ctx = nullptr;
    rc = ntc_ctx_create_for(device, idx, cmd == "encode" ? NTC_USE_ENCODE : NTC_USE_DECODE, &ctx);
    ntc_index_free(idx);
    if (rc) return fail("context", rc);
    uint64_t st[6] = {0};
    const double t1 = now();
    if (cmd == "encode") {
        fprintf(stderr, "Encoding fastX data...\n");
        rc = ntc_encode_file(ctx, inputs[0].c_str(), out.c_str(), threads, st);
        if (rc == 0)
            fprintf(stderr, "%llu reads, %llu bases -> %llu records, %llu bytes in %llu blocks (%llu dropped: empty column) | %.2f s (+%.2f s index)\n",
                    (unsigned long long)st[0], (unsigned long long)st[1], (unsigned long long)st[2], (unsigned long long)st[3],
                    (unsigned long long)st[4], (unsigned long long)st[5], now() - t1, t1 - t0);
    } else {
        fprintf(stderr, "Decoding encoded data...\n");
        rc = ntc_decode_file(ctx, inputs[0].c_str(), out.c_str(), threads, st);
        if (rc == 0)
            fprintf(stderr, "%llu blocks, %llu records -> %llu sequences, %llu bytes of FASTA | %.2f s (+%.2f s index)\n", (unsigned long long)st[4],
                    (unsigned long long)st[2], (unsigned long long)st[0], (unsigned long long)st[1], now() - t1, t1 - t0);
    }
    ntc_ctx_destroy(ctx);
    return rc ? fail(cmd.c_str(), rc) : 0;
}
