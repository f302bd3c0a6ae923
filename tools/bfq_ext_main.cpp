// bfq_ext_main.cpp -- drop-in for src_ext_mem/bfq_ext (the step-3 tool BFQzip_ext.py:199-220 runs):
// (OUT.bwt, OUT.bwt.qs, OUT.1.lcp) -> OUT.fq, clusters / smoothing / noise reduction / inverse
// eBWT on the GPU through the C ABI (include/bfq_gpu.h).  One binary per compile-time mode
// (src_ext_mem/makefile:10-22: -DM=0..3 -DB=0|1): it links the matching libbfqgpu_M<M>_B<B>.so.
//
// Same getopt string, defaults and "0 / -1 means default" rules as bfq_ext.cpp:969-1031; a
// missing required flag or input file prints the help and exits 0, as the reference does
// (bfq_ext.cpp:104-129, 1033-1054).  The statistics block on stdout has the reference's wording
// (bfq_ext.cpp:1113-1128) because BFQzip_ext.py captures it into OUT.log.  No ./tmp is needed.
#include "fastq_io.hpp"
#include "../include/bfq_gpu.h"
#include <stdlib.h>
#include <getopt.h>
#include <iostream>
#include <chrono>
#include <future>
#include <tuple>
#include <string>
#include <memory>
#include <algorithm>

using std::cout;
using std::endl;

static void help()
{
    cout << "bfq_ext [options]" << endl
         << "Options:" << endl
         << "-h          Print this help." << endl
         << "-e <arg>    Input eBWT file (A,C,G,T,#) of DNA (REQUIRED)." << endl
         << "-q <arg>    Qualities permuted according to the DNA's ebwt (REQUIRED)." << endl
         << "-a <arg>    LCP array file (REQUIRED)." << endl
         << "-o <arg>    Output fastq (REQUIRED)." << endl
         << "-l <arg>    Maximum read length (REQUIRED)." << endl
         << "-k <arg>    Minimum LCP required in clusters. Default: 16." << endl
         << "-m <arg>    Minimum length of cluster to be processed. Default: 2." << endl
         << "-v <arg>    Quality score for constant replacement (if M=2). Default: 29." << endl
         << "-f <arg>    Percentage threshold for frequent bases in clusters. Default: 40." << endl
         << "-t <arg>    Quality score threshold for trusted bases. Default: 20." << endl
         << "-s <arg>    ASCII value of end-marker symbol. Default: 35 (#)." << endl
         << "-T <arg>    Number of threads (ignored: the work runs on the GPU). Default: 1." << endl
         << "-G <arg>    CUDA device index. Default: 0." << endl
         << "-C          Also write OUT.fq.dna and OUT.fq.qs (the streams of the driver's step 4)." << endl
         << "-R          -C plus OUT.fq.qs.rlv / OUT.fq.qs.rll: run values and run lengths of OUT.fq.qs." << endl
         << "-H <arg>    Original headers from file." << endl
         << endl
         << "GPU build (sm_100a), mode M=" << bfq_mode_M() << " B=" << bfq_mode_B() << "." << endl;
    exit(0);
}

static bool file_exists(const std::string &p)
{
    struct stat st;
    return stat(p.c_str(), &st) == 0;
}

int main(int argc, char **argv)
{
    if (argc < 3) help();
    std::string input_dna, input_qual, input_lcp, output, input_titles;
    bool headers = false;
    int K = -1, m = 0, qthr = -1, maxlen = 0, device = 0, threads = 0;
    char default_value = '\0';
    float fthr = 0.0f;
    char TERM = '#';
    bool components = false, rle = false;
    int opt;
    while ((opt = getopt(argc, argv, "he:q:o:a:l:k:m:v:f:t:s:T:H:G:DVCR")) != -1) {
        switch (opt) {
        case 'h': help(); break;
        case 'e': input_dna = optarg; break;
        case 'q': input_qual = optarg; break;
        case 'o': output = optarg; break;
        case 'a': input_lcp = optarg; break;
        case 'l': maxlen = (uint8_t)atoi(optarg); break;        // dataTypelenSeq is uchar (parameters.h:68)
        case 'k': K = atoi(optarg); break;
        case 'm': m = atoi(optarg); break;
        case 'v': default_value = (char)atoi(optarg); break;
        case 't': qthr = atoi(optarg); break;
        case 'f': fthr = (float)atoi(optarg); break;
        case 's': TERM = (char)atoi(optarg); break;
        case 'T': threads = atoi(optarg); break;
        case 'G': device = atoi(optarg); break;
        case 'H': input_titles = optarg; headers = true; break;
        case 'D': case 'V': break;
        case 'C': components = true; break;                     // extension: also write OUT.fq.dna / OUT.fq.qs (step 4's streams)
        case 'R': components = true; rle = true; break;         // extension: -C plus the run-length form of OUT.fq.qs (.rlv values, .rll lengths)
        default: help(); return -1;
        }
    }
    (void)threads;
    bfq_params P;
    bfq_params_default(&P);
    if (K != -1) P.K = K;
    if (m != 0) P.m = m;
    if (default_value != '\0') P.default_q = (uint8_t)default_value;
    if (qthr != -1) P.qthr = qthr;
    if (fthr != 0.0f) P.fthr = fthr;
    P.eos_byte = (uint8_t)TERM;
    P.max_len = maxlen;
    if (input_dna.empty() || input_qual.empty() || output.empty() || input_lcp.empty() || maxlen == 0 || (headers && input_titles.empty())) help();
    for (const std::string *f : {&input_dna, &input_qual, &input_lcp}) {
        if (!file_exists(*f)) { cout << "Error: could not find file " << *f << "." << endl << endl; help(); }
    }
    if (headers && !file_exists(input_titles)) { cout << "Error: could not find file " << input_titles << "." << endl << endl; help(); }

    cout << "Running bfq_ext..." << endl << "\tMode: " << bfq_mode_M();
    if (bfq_mode_M() == 2) cout << "\treplacing QS with symbol: " << (char)P.default_q;
    cout << "\n\tK: " << P.K << endl << "\tm: " << P.m << endl << "\tFrequency threshold: " << P.fthr << "%" << endl << endl;
    cout << "Output fastq file: " << output << endl << endl;

    try {
        auto t0 = std::chrono::steady_clock::now();
        // the CUDA context comes up (seconds on a cold GPU) while the inputs are mapped and counted
        auto ctx_ready = std::async(std::launch::async, [device] {
            bfq_ctx *x = nullptr;
            int rc = bfq_ctx_create(&x, device);
            return std::make_tuple(rc, x, std::string(rc == BFQ_OK ? "" : bfq_last_error()));     // the message is thread-local
        });
        bfqio::MappedFile fb(input_dna), fq(input_qual), fl(input_lcp);
        const uint64_t n = fb.size;
        if (fq.size != n || fl.size != n) {
            std::cerr << "Error: " << input_dna << ", " << input_qual << " and " << input_lcp << " must have the same length (1-byte LCP entries)" << endl;
            bfq_ctx_destroy(std::get<1>(ctx_ready.get()));
            return EXIT_FAILURE;
        }
        uint64_t nreads = 0;
        for (uint64_t i = 0; i < n; i++) nreads += fb.data[i] == P.eos_byte;     // numSeq (bfq_ext.cpp:160-166)
        // Without a header file the records ("@" headers) are assembled on the device and arrive as the text of OUT.fq;
        // with -H the flat reads come back and the records are formatted here around the caller's header lines.
        const bool device_emit = (!headers && !getenv("BFQ_HOST_EMIT")) || components;
        if (components && headers) { std::cerr << "Error: -C (component streams) is not combined with -H" << endl; return EXIT_FAILURE; }
        const uint64_t text_cap = device_emit ? 2 * (n - nreads) + 6 * nreads + 64 : 1;
        const uint64_t flat_cap = device_emit ? 1 : (n ? n : 1);
        std::unique_ptr<uint8_t[]> text(new uint8_t[text_cap]), ob(new uint8_t[flat_cap]), oq(new uint8_t[flat_cap]);
        const uint64_t comp_cap = components ? n + 1 : 1;
        std::unique_ptr<uint8_t[]> dna(new uint8_t[comp_cap]), qsl(new uint8_t[comp_cap]);
        memset(dna.get(), 0, comp_cap); memset(qsl.get(), 0, comp_cap);
        std::vector<uint64_t> off(device_emit ? 1 : nreads + 1);
        uint64_t text_size = 0;
        // touch the output pages now (the device is still coming up): a download into untouched memory faults page by page
        memset(text.get(), 0, text_cap); memset(ob.get(), 0, flat_cap); memset(oq.get(), 0, flat_cap);
        auto tp = std::chrono::steady_clock::now();
        auto cr = ctx_ready.get();
        bfq_ctx *ctx = std::get<1>(cr);
        if (std::get<0>(cr) != BFQ_OK) { std::cerr << "Error: " << std::get<2>(cr) << endl; return EXIT_FAILURE; }
        bfq_stats S;
        auto t1 = std::chrono::steady_clock::now();
        int rc = device_emit ? bfq_smooth_rebuild_fastq(ctx, fb.data, fq.data, fl.data, n, &P, text.get(), text_cap, &text_size,
                                                        components ? dna.get() : nullptr, components ? qsl.get() : nullptr, &S)
                             : bfq_smooth_rebuild(ctx, fb.data, fq.data, fl.data, n, &P, ob.get(), oq.get(), off.data(), nreads + 1, &S);
        if (rc != BFQ_OK) { std::cerr << "Error: " << bfq_last_error() << endl; bfq_ctx_destroy(ctx); return EXIT_FAILURE; }
        // pre-transform for the host entropy coders, while the context is alive: run values and run
        // lengths of the quality stream (text = 2 T + 6 R bytes; the stream holds T + R)
        std::unique_ptr<uint8_t[]> rv, rl;
        uint64_t nruns = 0;
        if (rle) {
            const uint64_t k = (text_size - 6 * S.nreads) / 2 + S.nreads;
            rv.reset(new uint8_t[k + 1]); rl.reset(new uint8_t[k + 1]);
            if (bfq_qs_rle_encode(ctx, qsl.get(), k, rv.get(), rl.get(), k, &nruns) != BFQ_OK) {
                std::cerr << "Error: " << bfq_last_error() << endl; bfq_ctx_destroy(ctx); return EXIT_FAILURE;
            }
        }
        auto t2 = std::chrono::steady_clock::now();
        bfq_ctx_destroy(ctx);

        // ---- OUT.fq (decode.cpp:277-282: the output name is OUT + ".fq")
        std::vector<std::pair<uint64_t, uint32_t>> hidx;
        std::unique_ptr<bfqio::MappedFile> fh;
        if (headers) {
            fh.reset(new bfqio::MappedFile(input_titles));
            hidx = bfqio::index_lines(*fh);
            if (hidx.size() < nreads) { std::cerr << "Error: " << input_titles << " holds " << hidx.size() << " lines for " << nreads << " reads" << endl; return EXIT_FAILURE; }
        }
        FILE *fo = fopen((output + ".fq").c_str(), "wb");
        if (!fo) { std::cerr << "Error opening \"" << output << ".fq\" file" << endl; return EXIT_FAILURE; }
        if (device_emit) {
            if (text_size && fwrite(text.get(), 1, text_size, fo) != text_size) { std::cerr << "Error writing " << output << ".fq" << endl; return EXIT_FAILURE; }
        } else {
            std::vector<uint8_t> buf;
            const uint64_t CH = 1 << 16;
            for (uint64_t r0 = 0; r0 < nreads; r0 += CH) {
                buf.clear();
                bfqio::format_fastq(ob.get(), oq.get(), off.data(), r0, std::min(nreads, r0 + CH), headers ? fh->data : nullptr,
                                    headers ? hidx.data() : nullptr, buf);
                if (fwrite(buf.data(), 1, buf.size(), fo) != buf.size()) { std::cerr << "Error writing " << output << ".fq" << endl; return EXIT_FAILURE; }
            }
        }
        if (fclose(fo) != 0) { std::cerr << "Error closing " << output << ".fq" << endl; return EXIT_FAILURE; }
        if (components) {
            // text = 2 T + 6 R bytes with 4 R newlines... the streams hold T + R bytes: (text_size - 6 R) / 2 + R, R = S.nreads
            const uint64_t k = (text_size - 6 * S.nreads) / 2 + S.nreads;
            bfqio::write_all(output + ".fq.dna", dna.get(), k);
            bfqio::write_all(output + ".fq.qs", qsl.get(), k);
            if (rle) {
                bfqio::write_all(output + ".fq.qs.rlv", rv.get(), nruns);
                bfqio::write_all(output + ".fq.qs.rll", rl.get(), nruns);
            }
        }
        auto t3 = std::chrono::steady_clock::now();
        std::cerr << "\nThe rebuilt file is ready!\n\n";

        const uint64_t tot = S.clusters, nb = S.n - S.nreads;
        cout << "**** Cluster statistics ****" << endl;
        cout << tot << " is the total number of clusters" << endl;
        cout << "Discarded: " << S.discarded << "(" << (double(100 * S.discarded) / (tot)) << "%)" << endl;
        cout << "Ambiguous discarded: " << S.ambiguous << "(" << (double(100 * S.ambiguous) / (tot)) << "%)" << endl;
        cout << "Processed: " << S.processed << "(" << (double(100 * S.processed) / (tot)) << "%)" << endl << endl;
        cout << S.bases_inside << "/" << nb << " (" << (double(100 * S.bases_inside) / nb) << "%) bases fall inside clusters" << endl;
        cout << "**** Quality statistics ****" << endl;
        cout << S.qs_smoothed << "/" << nb << " qualities have been modified (" << 100 * double(S.qs_smoothed) / nb << "% of all qs and "
             << 100 * double(S.qs_smoothed) / S.bases_inside << "% of qs inside clusters)." << endl;
        cout << "***********************" << endl << endl;
        cout << "**** Bases statistics ****" << endl;
        cout << S.modified << "/" << nb << " bases have been modified (" << 100 * double(S.modified) / nb << "% of all bases and "
             << 100 * double(S.modified) / S.bases_inside << "% of bases inside clusters)." << endl;
        cout << "***********************" << endl << endl;
        auto sec = [](auto a, auto b) { return std::chrono::duration<double>(b - a).count(); };
        cout << "GPU timing: load " << sec(t0, t1) << " s (inputs " << sec(t0, tp) << " s), smooth+rebuild " << sec(t1, t2) << " s, write " << sec(t2, t3) << " s" << endl;
    } catch (std::exception &e) {
        std::cerr << "Error: " << e.what() << endl;
        return EXIT_FAILURE;
    }
    return 0;
}
