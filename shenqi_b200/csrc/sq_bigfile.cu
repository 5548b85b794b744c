// bigfile snapshot / IC reader: host code that reads the blocks the path needs - "<type>/Position",
// "<type>/Mass", "<type>/ID", the "Header" attributes - straight into the arrays sq_upload_particles_soa takes.
//
// Stands where the reference has petaio_read_snapshot / petaio_read_block (libgadget/petaio.cpp:908-1035 register the
// blocks, :590-700 read them) on top of the vendored bigfile library (depends/bigfile/src/bigfile.c).  The format, as
// that library writes it (bigfile.c:313-400, 587-620, 1600-1670):
//   <file>/<block>/header    text:  "DTYPE: <f8\nNMEMB: 3\nNFILE: 2\n" then per physical file "%06X: <rows> : <checksum> : <sysv sum>"
//   <file>/<block>/%06X      the rows of that physical file, raw, in the block's dtype (numpy-style: byte order, kind, size)
//   <file>/<block>/attr-v2   text, one attribute per line: "<name> <dtype> <nmemb> <hex of the raw bytes> #HUMANE [ ... ]"
// The per-file checksum is the byte sum modulo 2^32 (sysvsum, bigfile.c:1453-1460); it is verified on every file read
// in full.  Little-endian hosts only ('<', '=' or '|' byte order marks).  No MPI, no CUDA in this file.
#include "sq_common.cuh"
#include <errno.h>
#include <stdlib.h>
#include <string>
#include <vector>

namespace {
struct bf_block {
    std::string dir, dtype;
    int nmemb = 0, nfile = 0;
    std::vector<int64_t> rows;          // rows per physical file
    std::vector<unsigned> cksum;
    int64_t total = 0;
    int itemsize = 0;
    char kind = 0;
};

int bf_dtype(const std::string & dt, char & kind, int & size)
{
    size_t p = 0;
    if(p < dt.size() && (dt[p] == '<' || dt[p] == '=' || dt[p] == '|')) p++;
    else if(p < dt.size() && dt[p] == '>') return -1;
    if(p >= dt.size()) return -1;
    kind = dt[p++];
    size = atoi(dt.c_str() + p);
    if(!(kind == 'f' || kind == 'i' || kind == 'u' || kind == 'b' || kind == 'S' || kind == 'a') || size <= 0 || size > 16) return -1;
    return 0;
}

int bf_open(const char * path, const char * block, bf_block & B)
{
    B.dir = std::string(path) + "/" + block;
    const std::string hn = B.dir + "/header";
    FILE * f = fopen(hn.c_str(), "r");
    if(!f) { sq_set_error("bigfile: cannot open %s: %s", hn.c_str(), strerror(errno)); return SQ_ERR_ARG; }
    char dt[64] = "";
    int ok = fscanf(f, " DTYPE: %63s", dt) == 1 && fscanf(f, " NMEMB: %d", &B.nmemb) == 1 && fscanf(f, " NFILE: %d", &B.nfile) == 1;
    if(!ok || B.nmemb < 0 || B.nfile < 0 || B.nfile > (1 << 24)) { fclose(f); sq_set_error("bigfile: bad header %s", hn.c_str()); return SQ_ERR_ARG; }
    B.dtype = dt;
    if(bf_dtype(B.dtype, B.kind, B.itemsize)) { fclose(f); sq_set_error("bigfile: dtype %s of %s is not supported", dt, hn.c_str()); return SQ_ERR_ARG; }
    B.rows.assign(B.nfile, 0); B.cksum.assign(B.nfile, 0);
    B.total = 0;
    for(int i = 0; i < B.nfile; i++) {
        unsigned fid = 0, ck = 0, sysv = 0;
        long long sz = 0;
        if(fscanf(f, " %X: %lld : %u : %u", &fid, &sz, &ck, &sysv) != 4 || (int) fid >= B.nfile || sz < 0) {
            fclose(f); sq_set_error("bigfile: bad file table in %s", hn.c_str()); return SQ_ERR_ARG;
        }
        B.rows[fid] = sz; B.cksum[fid] = ck;
    }
    fclose(f);
    for(int i = 0; i < B.nfile; i++) B.total += B.rows[i];
    return SQ_OK;
}

// one value of the file's type -> double / int64 (exact for every integer the formats can hold up to 2^63)
template <class OUT>
void bf_convert(const unsigned char * src, char kind, int size, int64_t n, OUT * out)
{
    for(int64_t i = 0; i < n; i++) {
        const unsigned char * p = src + (size_t) i * size;
        if(kind == 'f' && size == 8) { double v; memcpy(&v, p, 8); out[i] = (OUT) v; }
        else if(kind == 'f' && size == 4) { float v; memcpy(&v, p, 4); out[i] = (OUT) v; }
        else if(kind == 'i' && size == 8) { int64_t v; memcpy(&v, p, 8); out[i] = (OUT) v; }
        else if(kind == 'i' && size == 4) { int32_t v; memcpy(&v, p, 4); out[i] = (OUT) v; }
        else if(kind == 'u' && size == 8) { uint64_t v; memcpy(&v, p, 8); out[i] = (OUT) v; }
        else if(kind == 'u' && size == 4) { uint32_t v; memcpy(&v, p, 4); out[i] = (OUT) v; }
        else if(size == 1) out[i] = (OUT) p[0];
        else if(size == 2 && kind == 'i') { int16_t v; memcpy(&v, p, 2); out[i] = (OUT) v; }
        else if(size == 2) { uint16_t v; memcpy(&v, p, 2); out[i] = (OUT) v; }
        else out[i] = (OUT) 0;
    }
}

int bf_out_size(const char * want)
{
    if(!strcmp(want, "f8") || !strcmp(want, "u8") || !strcmp(want, "i8")) return 8;
    if(!strcmp(want, "f4") || !strcmp(want, "i4") || !strcmp(want, "u4")) return 4;
    if(!strcmp(want, "u1")) return 1;
    return 0;
}
}

extern "C" {

/* dtype (at least 8 bytes), members per row, rows and physical files of a block */
int sq_bigfile_block_info(const char * path, const char * block, char * dtype, int * nmemb, int64_t * nrows, int * nfile)
{
    if(!path || !block) return SQ_ERR_ARG;
    bf_block B;
    SQ_TRY(bf_open(path, block, B));
    if(dtype) { strncpy(dtype, B.dtype.c_str(), 7); dtype[7] = 0; }
    if(nmemb) *nmemb = B.nmemb;
    if(nrows) *nrows = B.total;
    if(nfile) *nfile = B.nfile;
    return SQ_OK;
}

/* rows [first, first + count) of a block into out as `want` = "f8", "f4", "u8", "i8", "u4", "i4" or "u1" (count x nmemb
 * values, row-major); the checksum of every physical file that is read completely is verified */
int sq_bigfile_read_block(const char * path, const char * block, int64_t first, int64_t count, void * out, const char * want)
{
    if(!path || !block || !want || (count > 0 && !out) || first < 0 || count < 0) { sq_set_error("sq_bigfile_read_block: bad argument"); return SQ_ERR_ARG; }
    const int osz = bf_out_size(want);
    if(!osz) { sq_set_error("sq_bigfile_read_block: output type %s is not supported", want); return SQ_ERR_ARG; }
    bf_block B;
    SQ_TRY(bf_open(path, block, B));
    if(first + count > B.total) { sq_set_error("sq_bigfile_read_block: rows %lld..%lld of %lld in %s", (long long) first, (long long) (first + count), (long long) B.total, B.dir.c_str()); return SQ_ERR_ARG; }
    const size_t rowbytes = (size_t) B.itemsize * B.nmemb;
    int64_t fstart = 0, done = 0;
    std::vector<unsigned char> buf;
    for(int i = 0; i < B.nfile && done < count; i++) {
        const int64_t fend = fstart + B.rows[i];
        const int64_t a = first + done > fstart ? first + done : fstart, b = first + count < fend ? first + count : fend;
        if(b > a) {
            char name[16];
            snprintf(name, sizeof(name), "%06X", i);
            const std::string fn = B.dir + "/" + name;
            FILE * f = fopen(fn.c_str(), "rb");
            if(!f) { sq_set_error("bigfile: cannot open %s: %s", fn.c_str(), strerror(errno)); return SQ_ERR_ARG; }
            const bool whole = a == fstart && b == fend;
            const int64_t n = b - a;
            buf.resize((size_t) n * rowbytes + 1);
            if(fseek(f, (long) ((a - fstart) * (int64_t) rowbytes), SEEK_SET) || fread(buf.data(), rowbytes, (size_t) n, f) != (size_t) n) {
                fclose(f); sq_set_error("bigfile: short read in %s", fn.c_str()); return SQ_ERR_ARG;
            }
            fclose(f);
            if(whole) {
                unsigned sum = 0;
                for(size_t k = 0; k < (size_t) n * rowbytes; k++) sum += buf[k];
                if(sum != B.cksum[i]) { sq_set_error("bigfile: checksum of %s is %u, header says %u", fn.c_str(), sum, B.cksum[i]); return SQ_ERR_ARG; }
            }
            const int64_t nv = n * B.nmemb;
            unsigned char * o = (unsigned char *) out + (size_t) done * B.nmemb * osz;
            if(!strcmp(want, "f8")) bf_convert<double>(buf.data(), B.kind, B.itemsize, nv, (double *) o);
            else if(!strcmp(want, "f4")) bf_convert<float>(buf.data(), B.kind, B.itemsize, nv, (float *) o);
            else if(!strcmp(want, "u8")) bf_convert<uint64_t>(buf.data(), B.kind, B.itemsize, nv, (uint64_t *) o);
            else if(!strcmp(want, "i8")) bf_convert<int64_t>(buf.data(), B.kind, B.itemsize, nv, (int64_t *) o);
            else if(!strcmp(want, "u4")) bf_convert<uint32_t>(buf.data(), B.kind, B.itemsize, nv, (uint32_t *) o);
            else if(!strcmp(want, "i4")) bf_convert<int32_t>(buf.data(), B.kind, B.itemsize, nv, (int32_t *) o);
            else bf_convert<unsigned char>(buf.data(), B.kind, B.itemsize, nv, (unsigned char *) o);
            done += n;
        }
        fstart = fend;
    }
    return SQ_OK;
}

/* a numeric attribute of a block (attr-v2) as doubles: values[0 .. *n), at most maxn */
int sq_bigfile_read_attr(const char * path, const char * block, const char * name, double * values, int maxn, int * n)
{
    if(!path || !block || !name || !values || !n) return SQ_ERR_ARG;
    const std::string fn = std::string(path) + "/" + block + "/attr-v2";
    FILE * f = fopen(fn.c_str(), "r");
    if(!f) { sq_set_error("bigfile: cannot open %s: %s", fn.c_str(), strerror(errno)); return SQ_ERR_ARG; }
    std::vector<char> line(1 << 20);
    int rc = SQ_ERR_ARG;
    sq_set_error("bigfile: attribute %s not found in %s", name, fn.c_str());
    while(fgets(line.data(), (int) line.size(), f)) {
        char an[256], dt[32];
        int nm = 0, used = 0;
        if(sscanf(line.data(), "%255s %31s %d %n", an, dt, &nm, &used) < 3 || strcmp(an, name)) continue;
        char kind; int size;
        if(bf_dtype(dt, kind, size) || nm < 0) { sq_set_error("bigfile: attribute %s has dtype %s", name, dt); break; }
        const char * hex = line.data() + used;
        std::vector<unsigned char> raw((size_t) nm * size + 1);
        bool good = true;
        for(size_t k = 0; k < (size_t) nm * size; k++) {
            auto hv = [](char ch) { return ch >= '0' && ch <= '9' ? ch - '0' : (ch >= 'a' && ch <= 'f' ? ch - 'a' + 10 : (ch >= 'A' && ch <= 'F' ? ch - 'A' + 10 : -1)); };
            const int h = hv(hex[2 * k]), l = h < 0 ? -1 : hv(hex[2 * k + 1]);
            if(h < 0 || l < 0) { good = false; break; }
            raw[k] = (unsigned char) (h * 16 + l);
        }
        if(!good) { sq_set_error("bigfile: attribute %s: bad hex data", name); break; }
        const int m = nm < maxn ? nm : maxn;
        std::vector<double> v((size_t) nm + 1);
        bf_convert<double>(raw.data(), kind, size, nm, v.data());
        for(int k = 0; k < m; k++) values[k] = v[k];
        *n = nm;
        rc = SQ_OK;
        break;
    }
    fclose(f);
    return rc;
}

/* Particles of one type of a snapshot / IC file into the context: "<type>/Position" (-> f8), "<type>/Mass" if the block
 * exists, else Header.MassTable[type] (petaio.cpp: a zero MassTable entry means a Mass block), rows [first, first + count)
 * (count < 0: all).  Header.BoxSize is returned if wanted.  Then sq_upload_particles_soa. */
int sq_upload_particles_bigfile(sq_ctx * c, const char * path, int ptype, int64_t first, int64_t count, double * boxsize)
{
    if(!c || !path || ptype < 0 || ptype > 5) { sq_set_error("sq_upload_particles_bigfile: bad argument"); return SQ_ERR_ARG; }
    char blk[64];
    snprintf(blk, sizeof(blk), "%d/Position", ptype);
    int nmemb = 0; int64_t nrows = 0;
    SQ_TRY(sq_bigfile_block_info(path, blk, nullptr, &nmemb, &nrows, nullptr));
    if(nmemb != 3) { sq_set_error("sq_upload_particles_bigfile: %s has %d members per row", blk, nmemb); return SQ_ERR_ARG; }
    if(first < 0) first = 0;
    if(count < 0) count = nrows - first;
    if(first + count > nrows) { sq_set_error("sq_upload_particles_bigfile: rows %lld..%lld of %lld", (long long) first, (long long) (first + count), (long long) nrows); return SQ_ERR_ARG; }
    std::vector<double> pos((size_t) count * 3 + 1);
    std::vector<float> mass((size_t) count + 1);
    SQ_TRY(sq_bigfile_read_block(path, blk, first, count, pos.data(), "f8"));
    double mt[6] = {0, 0, 0, 0, 0, 0};
    int nmt = 0;
    const int have_mt = sq_bigfile_read_attr(path, "Header", "MassTable", mt, 6, &nmt) == SQ_OK && nmt > ptype && mt[ptype] > 0;
    if(have_mt) for(int64_t i = 0; i < count; i++) mass[i] = (float) mt[ptype];
    else {
        snprintf(blk, sizeof(blk), "%d/Mass", ptype);
        SQ_TRY(sq_bigfile_read_block(path, blk, first, count, mass.data(), "f4"));
    }
    if(boxsize) {
        int nb = 0;
        SQ_TRY(sq_bigfile_read_attr(path, "Header", "BoxSize", boxsize, 1, &nb));
    }
    return sq_upload_particles_soa(c, pos.data(), mass.data(), nullptr, nullptr, count);
}

} /* extern "C" */
