// heat_b200/csrc/ingest.cu -- native LightGCN-format reader (host code).
//
// Replaces the per-line Python parser of the reference's frontend (cf_cpu/cf/datasets.py:31-40:
// `for line in f: line.split() ...`, tens of seconds at the Amazon-Books size and infeasible at
// BASELINE.json configs[4]).  The file is read once, cut into one chunk per hardware thread at
// line boundaries, every chunk is parsed by its own thread, and the rows come back in FILE order
// as (row_users, row_indptr, items) -- the CSR the frontend builds its uint64 arrays from.
// Format: one line per user, `user item item ...`, fields separated by `separator` (and/or blanks),
// empty lines skipped; anything that is not a non-negative integer is an error (with its line number).
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include "heat_ctx.cuh"

namespace heat {
namespace {

struct Chunk {
    std::vector<int64_t> users, counts, items;
    uint64_t lines = 0;
    std::string error;
    uint64_t error_line = 0;      // line number inside the chunk (1-based)
};

void parse_chunk(const char *p, const char *end, char sep, Chunk *out)
{
    while (p < end) {
        const char *eol = static_cast<const char *>(memchr(p, '\n', (size_t)(end - p)));
        if (!eol) eol = end;
        ++out->lines;
        int64_t fields = 0;
        const char *q = p;
        while (q < eol) {
            const char ch = *q;
            if (ch == ' ' || ch == '\t' || ch == '\r' || ch == sep) { ++q; continue; }
            if (ch < '0' || ch > '9') {
                if (out->error.empty()) { out->error = std::string("unexpected character '") + ch + "'"; out->error_line = out->lines; }
                return;
            }
            int64_t v = 0;
            while (q < eol && *q >= '0' && *q <= '9') { v = v * 10 + (*q - '0'); ++q; }
            if (fields == 0) out->users.push_back(v); else out->items.push_back(v);
            ++fields;
        }
        if (fields > 0) out->counts.push_back(fields - 1);
        p = eol + 1;
    }
}

}  // namespace
}  // namespace heat

struct heat_clickfile {
    std::vector<heat::Chunk> chunks;
    uint64_t rows = 0, nnz = 0;
};

extern "C" {

int heat_clickfile_open(const char *path, char separator, heat_clickfile **out, uint64_t *num_rows, uint64_t *nnz)
{
    using namespace heat;
    if (!path || !out || !num_rows || !nnz) return fail("heat_clickfile_open: null argument");
    FILE *f = fopen(path, "rb");
    if (!f) return fail(std::string("heat_clickfile_open: cannot open ") + path);
    fseek(f, 0, SEEK_END);
    const long size = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<char> buf((size_t)std::max(size, 0L));
    const size_t got = buf.empty() ? 0 : fread(buf.data(), 1, buf.size(), f);
    fclose(f);
    if (got != buf.size()) return fail(std::string("heat_clickfile_open: short read of ") + path);
    const char *base = buf.data(), *end = base + buf.size();
    unsigned nt = std::max(1u, std::min(std::thread::hardware_concurrency(), 64u));
    if (buf.size() < (1u << 20)) nt = 1;
    std::vector<const char *> cut(nt + 1, end);
    cut[0] = base;
    for (unsigned t = 1; t < nt; ++t) {
        const char *p = base + buf.size() / nt * t;
        if (p < cut[t - 1]) p = cut[t - 1];
        const char *nl = static_cast<const char *>(memchr(p, '\n', (size_t)(end - p)));
        cut[t] = nl ? nl + 1 : end;
    }
    heat_clickfile *h = new heat_clickfile();
    h->chunks.resize(nt);
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nt; ++t) th.emplace_back(parse_chunk, cut[t], cut[t + 1], separator, &h->chunks[t]);
    parse_chunk(cut[0], cut[1], separator, &h->chunks[0]);
    for (auto &x : th) x.join();
    uint64_t line0 = 0;
    for (auto &c : h->chunks) {
        if (!c.error.empty()) {
            const std::string msg = std::string(path) + ":" + std::to_string(line0 + c.error_line) + ": " + c.error;
            delete h;
            return fail("heat_clickfile_open: " + msg);
        }
        line0 += c.lines;
        h->rows += c.users.size();
        h->nnz += c.items.size();
    }
    *out = h; *num_rows = h->rows; *nnz = h->nnz;
    return 0;
}

int heat_clickfile_read(heat_clickfile *h, int64_t *row_users, int64_t *row_indptr, int64_t *items)
{
    using namespace heat;
    if (!h || !row_users || !row_indptr || !items) return fail("heat_clickfile_read: null argument");
    uint64_t r = 0, n = 0;
    row_indptr[0] = 0;
    for (auto &c : h->chunks) {
        std::copy(c.users.begin(), c.users.end(), row_users + r);
        for (size_t i = 0; i < c.counts.size(); ++i) row_indptr[r + i + 1] = row_indptr[r + i] + c.counts[i];
        std::copy(c.items.begin(), c.items.end(), items + n);
        r += c.users.size(); n += c.items.size();
    }
    return 0;
}

int heat_clickfile_close(heat_clickfile *h)
{
    delete h;
    return 0;
}

}  // extern "C"
