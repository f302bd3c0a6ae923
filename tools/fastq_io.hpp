// fastq_io.hpp -- host side of the file contracts: FASTQ -> flat read arrays and back.
//
// Keeps what the reference's loaders keep of a FASTQ file -- line 2 (bases) and line 4 (qualities)
// of every 4-line record, in file order (external/gsufsort/lib/file.c:182-275) -- and writes
// records exactly as src_ext_mem/decode.cpp:1040-1127 does: header line ("@" alone unless a
// header file is given with -H), bases, "+", qualities.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include <stdexcept>
#include <sys/mman.h>
#include <sys/stat.h>
#include <fcntl.h>
#include <unistd.h>

namespace bfqio {

struct MappedFile {
    const uint8_t *data = nullptr;
    size_t size = 0;
    int fd = -1;
    explicit MappedFile(const std::string &path)
    {
        fd = open(path.c_str(), O_RDONLY);
        if (fd < 0) throw std::runtime_error("cannot open " + path);
        struct stat st;
        if (fstat(fd, &st) != 0) { close(fd); throw std::runtime_error("cannot stat " + path); }
        size = (size_t)st.st_size;
        if (size) {
            // MAP_POPULATE: the pages are faulted in here (while the CUDA context is still coming up in its own thread),
            // not one by one inside the upload
            void *p = mmap(nullptr, size, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);
            if (p == MAP_FAILED) { close(fd); throw std::runtime_error("cannot mmap " + path); }
            madvise(p, size, MADV_SEQUENTIAL);
            data = (const uint8_t *)p;
        }
    }
    ~MappedFile()
    {
        if (data) munmap((void *)data, size);
        if (fd >= 0) close(fd);
    }
    MappedFile(const MappedFile &) = delete;
    MappedFile &operator=(const MappedFile &) = delete;
};

struct Reads {
    std::vector<uint8_t> bases, quals;
    std::vector<uint64_t> off{0};           // nreads + 1
    std::vector<std::pair<uint64_t, uint32_t>> hdr;   // (file offset, length) of every header line
    uint64_t nreads() const { return off.size() - 1; }
};

// Appends the records of `path` to R.  Like gsufsort's loader (lib/file.c:182-275) a record is
// four lines; a trailing partial record is ignored.  Empty reads are kept out, as cat_all drops
// them (lib/utils.c:229-236).  keep_headers records where line 1 of every record lies.
inline void load_fastq(const MappedFile &f, Reads &R, bool keep_headers = false)
{
    const uint8_t *p = f.data, *end = f.data + f.size;
    R.bases.reserve(R.bases.size() + f.size / 2);
    R.quals.reserve(R.quals.size() + f.size / 2);
    while (p < end) {
        const uint8_t *l[5];
        l[0] = p;
        bool ok = true;
        for (int k = 1; k <= 4; k++) {
            const uint8_t *nl = (const uint8_t *)memchr(l[k - 1], '\n', end - l[k - 1]);
            if (!nl) {
                if (k == 4 && l[3] < end) { l[4] = end + 1; break; }   // last line without '\n'
                ok = false;
                break;
            }
            l[k] = nl + 1;
        }
        if (!ok) break;
        size_t blen = (size_t)(l[2] - 1 - l[1]), qlen = (size_t)(l[4] - 1 - l[3]);
        if (blen && l[1][blen - 1] == '\r') blen--;
        if (qlen && l[3][qlen - 1] == '\r') qlen--;
        if (blen != qlen) throw std::runtime_error("FASTQ record with different base and quality lengths");
        if (blen) {
            R.bases.insert(R.bases.end(), l[1], l[1] + blen);
            R.quals.insert(R.quals.end(), l[3], l[3] + qlen);
            R.off.push_back(R.bases.size());
            if (keep_headers) R.hdr.push_back({(uint64_t)(l[0] - f.data), (uint32_t)(l[1] - 1 - l[0])});
        }
        p = l[4] <= end ? l[4] : end;
    }
}

inline void write_all(const std::string &path, const void *buf, size_t n)
{
    FILE *f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot create " + path);
    if (n && fwrite(buf, 1, n, f) != n) { fclose(f); throw std::runtime_error("short write on " + path); }
    if (fclose(f) != 0) throw std::runtime_error("cannot close " + path);
}

inline std::vector<uint8_t> read_all(const std::string &path)
{
    MappedFile m(path);
    return std::vector<uint8_t>(m.data, m.data + m.size);
}

// header lines of a -H file (one per read, decode.cpp:1051-1076)
inline std::vector<std::pair<uint64_t, uint32_t>> index_lines(const MappedFile &f)
{
    std::vector<std::pair<uint64_t, uint32_t>> v;
    const uint8_t *p = f.data, *end = f.data + f.size;
    while (p < end) {
        const uint8_t *nl = (const uint8_t *)memchr(p, '\n', end - p);
        const uint8_t *e = nl ? nl : end;
        v.push_back({(uint64_t)(p - f.data), (uint32_t)(e - p)});
        p = e + 1;
    }
    return v;
}

// Formats reads [r0, r1) into `out` (appended).  hdr_src/hdr index the header lines (or null -> "@").
inline void format_fastq(const uint8_t *bases, const uint8_t *quals, const uint64_t *off, uint64_t r0, uint64_t r1,
                         const uint8_t *hdr_src, const std::pair<uint64_t, uint32_t> *hdr, std::vector<uint8_t> &out)
{
    size_t need = 0;
    for (uint64_t g = r0; g < r1; g++) need += 2 * (off[g + 1] - off[g]) + 5 + (hdr ? hdr[g].second : 1);
    size_t w = out.size();
    out.resize(w + need);
    uint8_t *o = out.data();
    for (uint64_t g = r0; g < r1; g++) {
        const size_t len = off[g + 1] - off[g];
        if (hdr) { memcpy(o + w, hdr_src + hdr[g].first, hdr[g].second); w += hdr[g].second; }
        else o[w++] = '@';
        o[w++] = '\n';
        memcpy(o + w, bases + off[g], len); w += len;
        o[w++] = '\n'; o[w++] = '+'; o[w++] = '\n';
        memcpy(o + w, quals + off[g], len); w += len;
        o[w++] = '\n';
    }
}

} // namespace bfqio
