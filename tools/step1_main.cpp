// step1_main.cpp -- drop-in for the step-1 tool of the BFQzip pipelines: FASTQ -> OUT.bwt,
// OUT.bwt.qs, OUT.<b>.lcp, computed on the GPU through the C ABI (include/bfq_gpu.h).
//
// One binary, two command-line spellings, chosen by the name it is invoked under:
//   eGap      IN.fastq --em --mem <MB> --qs -o OUT --lcp --lbytes 1     (BFQzip_ext.py:165-183;
//             options of external/egap/eGap:55-80).  End-marker byte 0, like eGap writes.
//   gsufsort  IN.fastq --bwt --qs --lcp 1 -o OUT                        (BFQzip.py:184; options
//             of external/gsufsort/main.c:113-146).  End-marker byte '#' (lib/utils.h:57-63).
// `--mem 0` is accepted (the driver passes it for inputs < 1 MiB).  Options that ask
// for structures outside this path (SA, DA, --rev, --bwt input ...) are refused with a non-zero
// exit status: the driver maps any non-zero status to failure (BFQzip_ext.py:310-318).
#include "fastq_io.hpp"
#include "../include/bfq_gpu.h"
#include <stdlib.h>
#include <chrono>
#include <future>
#include <memory>
#include <tuple>
#include <string>

static const char *base_name(const char *p)
{
    const char *s = strrchr(p, '/');
    return s ? s + 1 : p;
}

static void die(const std::string &m)
{
    fprintf(stderr, "Error: %s\n", m.c_str());
    exit(EXIT_FAILURE);
}

int main(int argc, char **argv)
{
    const bool gsuf_name = strstr(base_name(argv[0]), "gsufsort") != nullptr;
    bool gsuf = gsuf_name;
    std::vector<std::string> inputs;
    std::string out;
    bool want_lcp = false, want_qs = false, want_bwt = false;
    int lbytes = -1, device = 0, verbose = 0;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next_int = [&](int &dst) {
            if (i + 1 < argc && argv[i + 1][0] != '-') { dst = atoi(argv[++i]); return true; }
            return false;
        };
        if (a == "--style=egap") gsuf = false;
        else if (a == "--style=gsufsort") gsuf = true;
        else if (a == "-o" || a == "--out" || a == "--output") { if (++i >= argc) die("-o needs a value"); out = argv[i]; }
        else if (a == "--qs" || (a == "-q" && !gsuf) || (a == "-Q" && gsuf)) want_qs = true;
        else if (gsuf && (a == "-q" || a == "-l")) {}         // gsufsort: -q = FASTQ input, -l = light (main.c:135-137)
        else if (a == "--bwt" || a == "-B") { want_bwt = true; if (!gsuf) die("eGap --bwt (inputs are BWT files) is outside this path"); }
        else if (a == "--lcp" || (a == "-l" && !gsuf) || (a == "-L" && gsuf)) {
            want_lcp = true;
            if (gsuf) next_int(lbytes);                        // gsufsort: optional byte count (main.c:183-185)
        }
        else if (a == "--lbytes") { if (!next_int(lbytes)) die("--lbytes needs a value"); }
        else if (a == "--mem" || a == "-m") { int mem; if (!next_int(mem)) die("--mem needs a value"); (void)mem; }
        else if (a == "--em" || a == "--se" || a == "--im" || a == "--fastq" || a == "--build" || a == "--time") {}
        else if (a == "-v" || a == "--verbose") verbose++;
        else if (a == "--device") { if (!next_int(device)) die("--device needs a value"); }
        else if (a == "-h" || a == "--help") {
            printf("%s IN.fastq %s -o OUT\n  GPU drop-in (sm_100a): writes OUT.bwt, OUT.bwt.qs, OUT.<b>.lcp\n", argv[0],
                   gsuf ? "--bwt --qs --lcp 1" : "--em --mem M --qs --lcp --lbytes 1");
            return 0;
        }
        else if (a.size() > 1 && a[0] == '-') die("option " + a + " is outside the eBWT/QS/LCP path this tool replaces");
        else inputs.push_back(a);
    }
    if (inputs.empty()) die("no input file");
    if (gsuf && inputs.size() != 1) die("gsufsort takes exactly one input (main.c:232-235)");
    if (out.empty()) out = gsuf ? base_name(inputs[0].c_str()) : inputs[0];
    if (lbytes < 0) lbytes = gsuf ? 4 : 2;                     // tool defaults (main.c:92, eGap:66)
    if (lbytes != 1 && lbytes != 2 && lbytes != 4) die("LCP entries of 1, 2 or 4 bytes are supported");
    (void)want_bwt;

    try {
        auto t0 = std::chrono::steady_clock::now();
        // the CUDA context comes up (seconds on a cold GPU) while the FASTQ is parsed
        auto ctx_ready = std::async(std::launch::async, [device] {
            bfq_ctx *x = nullptr;
            int rc = bfq_ctx_create(&x, device);
            return std::make_tuple(rc, x, std::string(rc == BFQ_OK ? "" : bfq_last_error()));     // the message is thread-local
        });
        const uint8_t eos = gsuf ? (uint8_t)'#' : (uint8_t)0;
        // One input (what BFQzip_ext.py passes): the file goes to the device as it is and is parsed there.  Several inputs
        // (eGap concatenates them into one collection) are parsed on the host.
        const bool device_parse = inputs.size() == 1 && !getenv("BFQ_HOST_PARSE");
        std::unique_ptr<bfqio::MappedFile> text;
        bfqio::Reads R;
        uint64_t nreads = 0, N = 0, cap = 0;
        if (device_parse) {
            text.reset(new bfqio::MappedFile(inputs[0]));
            cap = text->size / 2 + 1;
        } else {
            for (auto &p : inputs) {
                bfqio::MappedFile f(p);
                bfqio::load_fastq(f, R);
            }
            nreads = R.nreads(); N = R.bases.size() + nreads; cap = N + 1;
        }
        std::unique_ptr<uint8_t[]> bwt_buf(new uint8_t[cap]), qs_buf(new uint8_t[cap]), lcp_buf(new uint8_t[cap]);
        // touch the output pages now (the device is still coming up): a download into untouched memory faults page by page
        memset(bwt_buf.get(), 0, cap); memset(qs_buf.get(), 0, cap); memset(lcp_buf.get(), 0, cap);
        struct View { uint8_t *p; uint8_t *data() const { return p; } uint8_t operator[](uint64_t i) const { return p[i]; } } bwt{bwt_buf.get()}, qs{qs_buf.get()}, lcp{lcp_buf.get()};
        auto tp = std::chrono::steady_clock::now();
        auto cr = ctx_ready.get();
        bfq_ctx *ctx = std::get<1>(cr);
        if (std::get<0>(cr) != BFQ_OK) die(std::get<2>(cr).c_str());
        auto t1 = std::chrono::steady_clock::now();
        int rc = device_parse
                     ? bfq_build_ebwt_fastq(ctx, text->data, text->size, eos, bwt.data(), qs.data(), lcp.data(), cap, &nreads, &N)
                     : bfq_build_ebwt(ctx, R.bases.data(), R.quals.data(), R.off.data(), nreads, eos, bwt.data(), qs.data(), lcp.data());
        if (rc != BFQ_OK) die(bfq_last_error());
        auto t2 = std::chrono::steady_clock::now();
        bfq_ctx_destroy(ctx);
        bfqio::write_all(out + ".bwt", bwt.data(), N);
        if (want_qs) bfqio::write_all(out + ".bwt.qs", qs.data(), N);
        if (want_lcp) {
            // LCP values fit one byte (reads <= 250 bp); wider entries are zero-extended, little-endian,
            // as both tools write the low lbytes of an int (main.c:891-895)
            const std::string name = out + "." + std::to_string(lbytes) + ".lcp";
            if (lbytes == 1) bfqio::write_all(name, lcp.data(), N);
            else {
                std::vector<uint8_t> wide(N * (size_t)lbytes, 0);
                for (uint64_t i = 0; i < N; i++) wide[i * lbytes] = lcp[i];
                bfqio::write_all(name, wide.data(), wide.size());
            }
        }
        auto t3 = std::chrono::steady_clock::now();
        auto sec = [](auto a, auto b) { return std::chrono::duration<double>(b - a).count(); };
        printf("%s (GPU): %llu reads, %llu symbols; load %.3f s (inputs %.3f s, device ready after %.3f s), eBWT+QS+LCP %.3f s, write %.3f s\n", gsuf ? "gsufsort" : "eGap",
               (unsigned long long)nreads, (unsigned long long)N, sec(t0, t1), sec(t0, tp), sec(t0, t1), sec(t1, t2), sec(t2, t3));
        (void)verbose;
    } catch (std::exception &e) {
        die(e.what());
    }
    return 0;
}
