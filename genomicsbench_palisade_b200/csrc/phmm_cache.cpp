// Binary cache of flat batches and the loader of GATK's regression-vector
// format (8f row 4).  Host-only.
//
// Cache file layout (little endian):
//   header, 128 bytes: magic "PHMMSOA1", u32 version, u32 n_sections (= 12), i64 n_batches,
//                      i64 n_reads, i64 n_haps, i64 read_pool_bytes, i64 hap_pool_bytes,
//                      u64 checksum of the section table, zero padding
//   section table    : n_sections x {u64 offset, u64 bytes, u64 fnv1a64}
//   sections, each 64-byte aligned, in the order of phmm_flat_batches:
//     batch_num_reads, batch_num_haps (i32), read_len (i32), read_off (i64), rs, q, i, d, c (u8),
//     hap_len (i32), hap_off (i64), hap (u8)
// Pools are stored packed (reads and haplotypes in order, offsets rebuilt), so a shard that points
// into larger pools writes only what it references.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "phmm_host.h"

phmm_parsed::~phmm_parsed()
{
    if (map) munmap(map, map_len);
}

namespace {

const char MAGIC[8] = {'P', 'H', 'M', 'M', 'S', 'O', 'A', '1'};
const uint32_t VERSION = 1;
const int N_SECTIONS = 12;

struct Header {
    char magic[8];
    uint32_t version, n_sections;
    int64_t n_batches, n_reads, n_haps, read_pool_bytes, hap_pool_bytes;
    uint64_t table_sum;
    uint8_t pad[128 - 8 - 8 - 5 * 8 - 8];
};
static_assert(sizeof(Header) == 128, "header is 128 bytes");

struct Section { uint64_t off, bytes, sum; };

uint64_t fnv1a64(const void *p, size_t n)
{
    // 8 bytes at a time (the cache is read at memory speed; a byte-serial FNV would dominate open)
    const uint8_t *b = (const uint8_t *)p;
    uint64_t h = 0xcbf29ce484222325ull;
    size_t k = 0;
    for (; k + 8 <= n; k += 8) {
        uint64_t w;
        memcpy(&w, b + k, 8);
        h = (h ^ w) * 0x100000001b3ull;
    }
    for (; k < n; ++k) h = (h ^ b[k]) * 0x100000001b3ull;
    return h;
}

int fail(char *err, size_t err_len, int code, const char *fmt, const char *arg)
{
    if (err && err_len) snprintf(err, err_len, fmt, arg);
    return code;
}

}  // namespace

extern "C" int phmm_cache_write(const phmm_flat_batches *fb, const char *path, char *err, size_t err_len)
{
    if (!fb || !path) return PHMM_ERR_ARG;
    // pack: reads and haplotypes in order, pools rebuilt back to back
    phmm::ParsedBatches b;
    b.batch_num_reads.assign(fb->batch_num_reads, fb->batch_num_reads + fb->n_batches);
    b.batch_num_haps.assign(fb->batch_num_haps, fb->batch_num_haps + fb->n_batches);
    int64_t sr = 0, sh = 0;
    for (int64_t k = 0; k < fb->n_batches; ++k) { sr += fb->batch_num_reads[k]; sh += fb->batch_num_haps[k]; }
    if (sr != fb->n_reads || sh != fb->n_haps)
        return fail(err, err_len, PHMM_ERR_ARG, "%s: batch tables do not add up to n_reads / n_haps", path);
    for (int64_t r = 0; r < fb->n_reads; ++r) {
        const int32_t n = fb->read_len[r];
        const int64_t o = fb->read_off[r];
        if (n <= 0 || o < 0) return fail(err, err_len, PHMM_ERR_ARG, "%s: bad read length / offset", path);
        b.read_len.push_back(n);
        b.read_off.push_back((int64_t)b.rs.size());
        b.rs.insert(b.rs.end(), fb->rs + o, fb->rs + o + n);
        b.q.insert(b.q.end(), fb->q + o, fb->q + o + n);
        b.i.insert(b.i.end(), fb->i + o, fb->i + o + n);
        b.d.insert(b.d.end(), fb->d + o, fb->d + o + n);
        b.c.insert(b.c.end(), fb->c + o, fb->c + o + n);
    }
    for (int64_t h = 0; h < fb->n_haps; ++h) {
        const int32_t n = fb->hap_len[h];
        const int64_t o = fb->hap_off[h];
        if (n <= 0 || o < 0) return fail(err, err_len, PHMM_ERR_ARG, "%s: bad haplotype length / offset", path);
        b.hap_len.push_back(n);
        b.hap_off.push_back((int64_t)b.hap.size());
        b.hap.insert(b.hap.end(), fb->hap + o, fb->hap + o + n);
    }
    const struct { const void *p; size_t bytes; } src[N_SECTIONS] = {
        {b.batch_num_reads.data(), b.batch_num_reads.size() * 4}, {b.batch_num_haps.data(), b.batch_num_haps.size() * 4},
        {b.read_len.data(), b.read_len.size() * 4}, {b.read_off.data(), b.read_off.size() * 8},
        {b.rs.data(), b.rs.size()}, {b.q.data(), b.q.size()}, {b.i.data(), b.i.size()},
        {b.d.data(), b.d.size()}, {b.c.data(), b.c.size()},
        {b.hap_len.data(), b.hap_len.size() * 4}, {b.hap_off.data(), b.hap_off.size() * 8},
        {b.hap.data(), b.hap.size()},
    };
    Section tab[N_SECTIONS];
    uint64_t off = sizeof(Header) + sizeof tab;
    for (int k = 0; k < N_SECTIONS; ++k) {
        off = (off + 63) & ~(uint64_t)63;
        tab[k].off = off;
        tab[k].bytes = src[k].bytes;
        tab[k].sum = fnv1a64(src[k].p, src[k].bytes);
        off += src[k].bytes;
    }
    Header hd;
    memset(&hd, 0, sizeof hd);
    memcpy(hd.magic, MAGIC, 8);
    hd.version = VERSION;
    hd.n_sections = N_SECTIONS;
    hd.n_batches = fb->n_batches; hd.n_reads = fb->n_reads; hd.n_haps = fb->n_haps;
    hd.read_pool_bytes = (int64_t)b.rs.size(); hd.hap_pool_bytes = (int64_t)b.hap.size();
    hd.table_sum = fnv1a64(tab, sizeof tab);

    // write to a temporary name and rename, so a reader never sees a half-written cache
    std::string tmp = std::string(path) + ".tmp";
    FILE *f = fopen(tmp.c_str(), "wb");
    if (!f) return fail(err, err_len, PHMM_ERR_IO, "cannot create %s", tmp.c_str());
    bool ok = fwrite(&hd, sizeof hd, 1, f) == 1 && fwrite(tab, sizeof tab, 1, f) == 1;
    uint64_t at = sizeof hd + sizeof tab;
    static const uint8_t zeros[64] = {0};
    for (int k = 0; ok && k < N_SECTIONS; ++k) {
        if (tab[k].off > at) { ok = fwrite(zeros, (size_t)(tab[k].off - at), 1, f) == 1; at = tab[k].off; }
        if (ok && tab[k].bytes) ok = fwrite(src[k].p, tab[k].bytes, 1, f) == 1;
        at += tab[k].bytes;
    }
    ok = (fclose(f) == 0) && ok;
    if (!ok || rename(tmp.c_str(), path) != 0) {
        unlink(tmp.c_str());
        return fail(err, err_len, PHMM_ERR_IO, "cannot write %s", path);
    }
    return PHMM_OK;
}

extern "C" int phmm_cache_open(const char *path, phmm_parsed **out, char *err, size_t err_len)
{
    if (!path || !out) return PHMM_ERR_ARG;
    *out = nullptr;
    int fd = open(path, O_RDONLY);
    if (fd < 0) return fail(err, err_len, PHMM_ERR_IO, "cannot open %s", path);
    struct stat st;
    if (fstat(fd, &st) != 0) { close(fd); return fail(err, err_len, PHMM_ERR_IO, "cannot stat %s", path); }
    const size_t len = (size_t)st.st_size;
    if (len < sizeof(Header) + N_SECTIONS * sizeof(Section)) {
        close(fd);
        return fail(err, err_len, PHMM_ERR_PARSE, "%s: too short for a phmm cache", path);
    }
    void *m = mmap(nullptr, len, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (m == MAP_FAILED) return fail(err, err_len, PHMM_ERR_IO, "cannot mmap %s", path);
    phmm_parsed *p = new phmm_parsed();
    p->map = m;
    p->map_len = len;
    auto bad = [&](const char *what) {
        delete p;
        if (err && err_len) snprintf(err, err_len, "%s: %s", path, what);
        return PHMM_ERR_PARSE;
    };
    const uint8_t *base = (const uint8_t *)m;
    Header hd;
    memcpy(&hd, base, sizeof hd);
    if (memcmp(hd.magic, MAGIC, 8) != 0) return bad("not a phmm cache (magic)");
    if (hd.version != VERSION || hd.n_sections != (uint32_t)N_SECTIONS) return bad("unsupported cache version");
    Section tab[N_SECTIONS];
    memcpy(tab, base + sizeof hd, sizeof tab);
    if (fnv1a64(tab, sizeof tab) != hd.table_sum) return bad("section table checksum mismatch");
    if (hd.n_batches < 0 || hd.n_reads < 0 || hd.n_haps < 0 || hd.read_pool_bytes < 0 || hd.hap_pool_bytes < 0)
        return bad("negative count");
    // counts are bounded by the file length BEFORE they are multiplied: a crafted n_reads = 2^61 would wrap to a
    // small section size, pass the bounds test below and send the table walks far past the mapping
    if ((uint64_t)hd.n_batches > len / 4 || (uint64_t)hd.n_reads > len / 8 || (uint64_t)hd.n_haps > len / 8 ||
        (uint64_t)hd.read_pool_bytes > len || (uint64_t)hd.hap_pool_bytes > len)
        return bad("count larger than the file");
    const uint64_t want[N_SECTIONS] = {
        (uint64_t)hd.n_batches * 4, (uint64_t)hd.n_batches * 4, (uint64_t)hd.n_reads * 4, (uint64_t)hd.n_reads * 8,
        (uint64_t)hd.read_pool_bytes, (uint64_t)hd.read_pool_bytes, (uint64_t)hd.read_pool_bytes,
        (uint64_t)hd.read_pool_bytes, (uint64_t)hd.read_pool_bytes,
        (uint64_t)hd.n_haps * 4, (uint64_t)hd.n_haps * 8, (uint64_t)hd.hap_pool_bytes};
    for (int k = 0; k < N_SECTIONS; ++k) {
        if (tab[k].bytes != want[k] || (tab[k].off & 63) || tab[k].off > len || tab[k].bytes > len - tab[k].off)
            return bad("section out of bounds");
        if (fnv1a64(base + tab[k].off, tab[k].bytes) != tab[k].sum) return bad("section checksum mismatch (truncated or corrupted)");
    }
    phmm_flat_batches &v = p->b.view;
    memset(&v, 0, sizeof v);
    v.n_batches = hd.n_batches; v.n_reads = hd.n_reads; v.n_haps = hd.n_haps;
    v.batch_num_reads = (const int32_t *)(base + tab[0].off);
    v.batch_num_haps = (const int32_t *)(base + tab[1].off);
    v.read_len = (const int32_t *)(base + tab[2].off);
    v.read_off = (const int64_t *)(base + tab[3].off);
    v.rs = base + tab[4].off; v.q = base + tab[5].off; v.i = base + tab[6].off;
    v.d = base + tab[7].off; v.c = base + tab[8].off;
    v.hap_len = (const int32_t *)(base + tab[9].off);
    v.hap_off = (const int64_t *)(base + tab[10].off);
    v.hap = base + tab[11].off;
    // the tables must be consistent with the pools they index (a valid checksum only proves the
    // file is what was written)
    int64_t sr = 0, sh = 0;
    for (int64_t k = 0; k < v.n_batches; ++k) {
        if (v.batch_num_reads[k] < 0 || v.batch_num_haps[k] < 0) return bad("negative batch size");
        sr += v.batch_num_reads[k]; sh += v.batch_num_haps[k];
    }
    if (sr != v.n_reads || sh != v.n_haps) return bad("batch tables do not add up");
    for (int64_t r = 0; r < v.n_reads; ++r)
        if (v.read_len[r] <= 0 || v.read_off[r] < 0 || v.read_len[r] > hd.read_pool_bytes ||
            v.read_off[r] > hd.read_pool_bytes - v.read_len[r])
            return bad("read outside its pool");
    for (int64_t h = 0; h < v.n_haps; ++h)
        if (v.hap_len[h] <= 0 || v.hap_off[h] < 0 || v.hap_len[h] > hd.hap_pool_bytes ||
            v.hap_off[h] > hd.hap_pool_bytes - v.hap_len[h])
            return bad("haplotype outside its pool");
    *out = p;
    return PHMM_OK;
}

// ============================================================================================
// GATK regression vectors: <hap> <read> <quals> <ins> <del> <gcp> <expected>, '#' comments
// (PairHmmUnitTest.java:166-185; normalize :268-278)
// ============================================================================================
extern "C" int phmm_testdata_parse_buffer(const char *data, size_t len, phmm_parsed **out, char *err, size_t err_len)
{
    if (!out || (len && !data)) return PHMM_ERR_ARG;
    *out = nullptr;
    phmm_parsed *p = new phmm_parsed();
    phmm::ParsedBatches &b = p->b;
    long line = 1;
    auto bad = [&](const char *what) {
        delete p;
        if (err && err_len) snprintf(err, err_len, "line %ld: %s", line, what);
        return PHMM_ERR_PARSE;
    };
    auto is_ws = [](char ch) { return ch == ' ' || ch == '\n' || ch == '\t' || ch == '\r' || ch == '\v' || ch == '\f'; };
    const char *s = data, *end = data + len;
    auto norm = [](std::vector<uint8_t> &pool, const char *t, size_t n, int floor_v) {
        for (size_t k = 0; k < n; ++k) {
            const int v = (int)(signed char)((signed char)t[k] - 33);     // Java byte arithmetic
            pool.push_back((uint8_t)(v < floor_v ? floor_v : v));
        }
    };
    for (;;) {
        while (s < end && is_ws(*s)) { if (*s == '\n') ++line; ++s; }
        if (s >= end) break;
        if (*s == '#') {                       // Scanner.hasNext("#.*") -> nextLine()
            while (s < end && *s != '\n') ++s;
            continue;
        }
        const char *tok[7]; size_t n[7];
        for (int k = 0; k < 7; ++k) {
            while (s < end && is_ws(*s)) { if (*s == '\n') ++line; ++s; }
            if (s >= end) return bad("unexpected end of input inside a test vector");
            tok[k] = s;
            while (s < end && !is_ws(*s)) ++s;
            n[k] = (size_t)(s - tok[k]);
        }
        if (n[0] == 0 || n[0] > (size_t)PHMM_MAX_SEQ_LEN || n[1] == 0 || n[1] > (size_t)PHMM_MAX_SEQ_LEN)
            return bad("sequence length out of range");
        for (int k = 2; k < 6; ++k)
            if (n[k] != n[1]) return bad("quality string length differs from the read length");
        char num[64];
        if (n[6] >= sizeof num) return bad("expected value too long");
        memcpy(num, tok[6], n[6]);
        num[n[6]] = 0;
        char *stop = nullptr;
        errno = 0;
        const double e = strtod(num, &stop);
        if (stop == num || *stop != 0) return bad("expected value is not a number");
        b.batch_num_reads.push_back(1);
        b.batch_num_haps.push_back(1);
        b.hap_len.push_back((int32_t)n[0]);
        b.hap_off.push_back((int64_t)b.hap.size());
        b.hap.insert(b.hap.end(), (const uint8_t *)tok[0], (const uint8_t *)tok[0] + n[0]);
        b.read_len.push_back((int32_t)n[1]);
        b.read_off.push_back((int64_t)b.rs.size());
        b.rs.insert(b.rs.end(), (const uint8_t *)tok[1], (const uint8_t *)tok[1] + n[1]);
        norm(b.q, tok[2], n[2], 6);
        norm(b.i, tok[3], n[3], 0);
        norm(b.d, tok[4], n[4], 0);
        norm(b.c, tok[5], n[5], 0);
        p->expected.push_back(e);
    }
    b.finish();
    *out = p;
    return PHMM_OK;
}

extern "C" int phmm_testdata_parse_file(const char *path, phmm_parsed **out, char *err, size_t err_len)
{
    if (!path || !out) return PHMM_ERR_ARG;
    *out = nullptr;
    FILE *f = fopen(path, "rb");
    if (!f) return fail(err, err_len, PHMM_ERR_IO, "cannot open %s", path);
    std::string buf;
    char tmp[1 << 16];
    size_t n;
    while ((n = fread(tmp, 1, sizeof tmp, f)) > 0) buf.append(tmp, n);
    fclose(f);
    return phmm_testdata_parse_buffer(buf.data(), buf.size(), out, err, err_len);
}

extern "C" int64_t phmm_testdata_expected(const phmm_parsed *p, double *expected, int64_t cap)
{
    if (!p) return -1;
    const int64_t n = (int64_t)p->expected.size();
    if (expected && cap > 0) memcpy(expected, p->expected.data(), (size_t)std::min(n, cap) * sizeof(double));
    return n;
}
