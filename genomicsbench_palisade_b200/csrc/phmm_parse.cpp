// Text input of the phmm benchmark -> flat batches.
//
// Format, as read by read_batch / read_testfile (benchmarks/phmm/PairHMMUnitTest_orig.cpp:97-213,
// identical in tools/GKL/src/examples/PairHMM/PairHMMUnitTest.cpp:61-106): repeated to EOF
//     <num_reads> <num_haps>
//     num_reads lines:  <bases> <q> <i> <d> <c>      (whitespace separated, Phred+33)
//     num_haps lines:   <bases>
// The reference tokenises with operator>> (any whitespace) and normalises each quality string
// with max(floor, byte - 33), floor 6 for q and 0 for i/d/c (:88-92,:120-123).  It strndup()s
// the four quality strings to bases.size() without checking their length (:142-151); here a
// length mismatch is a parse error.  An empty file yields zero batches (the reference reads
// uninitialised counts in that case, :100).
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdio>
#include <cstring>

#include "phmm_host.h"

namespace phmm {

void ParsedBatches::finish()
{
    memset(&view, 0, sizeof view);
    view.n_batches = (int64_t)batch_num_reads.size();
    view.batch_num_reads = batch_num_reads.data();
    view.batch_num_haps = batch_num_haps.data();
    view.n_reads = (int64_t)read_len.size();
    view.read_len = read_len.data();
    view.read_off = read_off.data();
    // non-NULL pool addresses even when empty
    static const uint8_t dummy = 0;
    view.rs = rs.empty() ? &dummy : rs.data();
    view.q = q.empty() ? &dummy : q.data();
    view.i = i.empty() ? &dummy : i.data();
    view.d = d.empty() ? &dummy : d.data();
    view.c = c.empty() ? &dummy : c.data();
    view.n_haps = (int64_t)hap_len.size();
    view.hap_len = hap_len.data();
    view.hap_off = hap_off.data();
    view.hap = hap.empty() ? &dummy : hap.data();
}

namespace {

struct Cursor {
    const char *p, *end;
    long line = 1;
    static bool is_ws(char ch) { return ch == ' ' || ch == '\n' || ch == '\t' || ch == '\r' || ch == '\v' || ch == '\f'; }
    void skip_ws()
    {
        while (p < end && is_ws(*p)) { if (*p == '\n') ++line; ++p; }
    }
    bool token(const char **tok, size_t *len)
    {
        skip_ws();
        if (p >= end) return false;
        const char *s = p;
        while (p < end && !is_ws(*p)) ++p;
        *tok = s;
        *len = (size_t)(p - s);
        return true;
    }
    bool integer(long *v)
    {
        const char *t; size_t n;
        if (!token(&t, &n) || n == 0 || n > 10) return false;
        long x = 0;
        for (size_t k = 0; k < n; ++k) {
            if (t[k] < '0' || t[k] > '9') return false;
            x = x * 10 + (t[k] - '0');
        }
        *v = x;
        return true;
    }
};

void append_norm(std::vector<uint8_t> &pool, const char *s, size_t n, int floor_v)
{
    const size_t at = pool.size();
    pool.resize(at + n);
    for (size_t k = 0; k < n; ++k) {
        // the reference works on (signed) char: max(floor, s[k] - 33)
        const int v = (int)(signed char)s[k] - 33;
        pool[at + k] = (uint8_t)(v < floor_v ? floor_v : v);
    }
}

int parse(const char *data, size_t len, ParsedBatches *b, char *err, size_t err_len)
{
    Cursor cur{data, data + len};
    auto fail = [&](const char *what) {
        if (err && err_len) snprintf(err, err_len, "line %ld: %s", cur.line, what);
        return PHMM_ERR_PARSE;
    };
    b->rs.reserve(len / 6);
    b->q.reserve(len / 6); b->i.reserve(len / 6); b->d.reserve(len / 6); b->c.reserve(len / 6);
    for (;;) {
        cur.skip_ws();
        if (cur.p >= cur.end) break;
        long nr, nh;
        if (!cur.integer(&nr) || !cur.integer(&nh)) return fail("expected '<num_reads> <num_haps>'");
        if (nr > INT32_MAX || nh > INT32_MAX) return fail("batch too large");
        b->batch_num_reads.push_back((int32_t)nr);
        b->batch_num_haps.push_back((int32_t)nh);
        for (long r = 0; r < nr; ++r) {
            const char *f[5]; size_t n[5];
            for (int k = 0; k < 5; ++k)
                if (!cur.token(&f[k], &n[k])) return fail("unexpected end of input inside a read");
            if (n[0] == 0 || n[0] > (size_t)PHMM_MAX_SEQ_LEN) return fail("read length out of range");
            for (int k = 1; k < 5; ++k)
                if (n[k] != n[0]) return fail("quality string length differs from the read length");
            b->read_len.push_back((int32_t)n[0]);
            b->read_off.push_back((int64_t)b->rs.size());
            b->rs.insert(b->rs.end(), (const uint8_t *)f[0], (const uint8_t *)f[0] + n[0]);
            append_norm(b->q, f[1], n[1], 6);
            append_norm(b->i, f[2], n[2], 0);
            append_norm(b->d, f[3], n[3], 0);
            append_norm(b->c, f[4], n[4], 0);
        }
        for (long h = 0; h < nh; ++h) {
            const char *t; size_t n;
            if (!cur.token(&t, &n)) return fail("unexpected end of input inside the haplotypes");
            if (n == 0 || n > (size_t)PHMM_MAX_SEQ_LEN) return fail("haplotype length out of range");
            b->hap_len.push_back((int32_t)n);
            b->hap_off.push_back((int64_t)b->hap.size());
            b->hap.insert(b->hap.end(), (const uint8_t *)t, (const uint8_t *)t + n);
        }
    }
    b->finish();
    return PHMM_OK;
}

}  // namespace
}  // namespace phmm

extern "C" int phmm_parse_buffer(const char *data, size_t len, phmm_parsed **out, char *err, size_t err_len)
{
    if (!out || (len && !data)) return PHMM_ERR_ARG;
    *out = nullptr;
    phmm_parsed *p = new phmm_parsed();
    int rc = phmm::parse(data, len, &p->b, err, err_len);
    if (rc != PHMM_OK) { delete p; return rc; }
    *out = p;
    return PHMM_OK;
}

extern "C" int phmm_parse_file(const char *path, phmm_parsed **out, char *err, size_t err_len)
{
    if (!path || !out) return PHMM_ERR_ARG;
    *out = nullptr;
    int fd = open(path, O_RDONLY);
    if (fd < 0) { if (err && err_len) snprintf(err, err_len, "cannot open %s", path); return PHMM_ERR_IO; }
    struct stat st;
    if (fstat(fd, &st) != 0) { close(fd); if (err && err_len) snprintf(err, err_len, "cannot stat %s", path); return PHMM_ERR_IO; }
    int rc;
    if (st.st_size == 0) {
        rc = phmm_parse_buffer("", 0, out, err, err_len);
    } else {
        void *m = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) { close(fd); if (err && err_len) snprintf(err, err_len, "cannot mmap %s", path); return PHMM_ERR_IO; }
        madvise(m, (size_t)st.st_size, MADV_SEQUENTIAL);
        rc = phmm_parse_buffer((const char *)m, (size_t)st.st_size, out, err, err_len);
        munmap(m, (size_t)st.st_size);
    }
    close(fd);
    return rc;
}

extern "C" const phmm_flat_batches *phmm_parsed_batches(const phmm_parsed *p) { return p ? &p->b.view : nullptr; }
extern "C" void phmm_parsed_free(phmm_parsed *p) { delete p; }
