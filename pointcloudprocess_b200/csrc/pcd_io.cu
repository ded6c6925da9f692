// pcd_io.cu -- host side of the path's data format: PCD v0.7 files with ascii, binary and
// binary_compressed payloads, and the LZF codec the compressed payload uses.  Replaces io::loadPCDFile /
// savePCDFile* (pcd_helper.h:1343-1497), PCDReader::readHeader / read (pcd_helper.cpp:71-333, 845-1181),
// PCDWriter::generateHeader / writeASCII / writeBinary / writeBinaryCompressed (pcd_helper.h:323-371, 489-1090)
// and lzfCompress / lzfDecompress (lzf.cpp:86-415).  Files are read into / written from the caller's PointXYZRGBA
// (or xyz-only) records on the host; pcp_pcd_load_many is the pipelined loader: payloads are read and
// LZF-decoded by host threads, uploaded AS STORED (packed records or field planes, 32 B/point instead of 48) and the
// records are assembled by a kernel.
#include "common.cuh"
#include <cerrno>
#include <cstdarg>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <future>
#include <memory>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

namespace {

// ---- LZF (liblzf format, HLOG 13, the reference's default parse) -------------------------------------------
//   000LLLLL <L+1 bytes>        literal run
//   LLLooooo oooooooo           match of L+2 bytes (L = 1..6) at distance o+1
//   111ooooo LLLLLLLL oooooooo  match of L+9 bytes at distance o+1
struct LzfWriter {
    unsigned char* out;
    size_t cap, at = 1, run = 0; // at = 1: byte 0 is the header of the first literal run
    bool ok = true;
    LzfWriter(unsigned char* o, size_t c) : out(o), cap(c) {}
    void close_run() // the header of the open literal run, or nothing if the run is empty
    {
        out[at - run - 1] = (unsigned char)(run - 1);
        if (!run) at--;
    }
    void literal(unsigned char b)
    {
        if (at >= cap) {
            ok = false;
            return;
        }
        out[at++] = b;
        if (++run == 32) {
            out[at - run - 1] = 31;
            run = 0;
            at++;
        }
    }
    bool room_for_match() const { return !(at + 4 >= cap && at - (run ? 0 : 1) + 4 >= cap); }
    void match(size_t dist_m1, size_t len_m2)
    {
        close_run();
        if (len_m2 < 7) {
            out[at++] = (unsigned char)((dist_m1 >> 8) + (len_m2 << 5));
        } else {
            out[at++] = (unsigned char)((dist_m1 >> 8) + (7u << 5));
            out[at++] = (unsigned char)(len_m2 - 7);
        }
        out[at++] = (unsigned char)dist_m1;
        run = 0;
        at++; // header of the next run
    }
};

inline unsigned lzf_slot(unsigned h3) { return ((h3 >> (24 - 13)) - h3) & 0x1fffu; }
inline unsigned lzf_h3(const unsigned char* p) { return ((unsigned)p[0] << 16) | ((unsigned)p[1] << 8) | p[2]; }

unsigned lzf_encode(const unsigned char* in, size_t n, unsigned char* out, size_t cap)
{
    if (!n || !cap) return 0;
    std::vector<unsigned> last(1u << 13, 0u); // position of the latest occurrence of each 3-byte hash
    LzfWriter w(out, cap);
    size_t i = 0;
    // rolling 3-byte window kept as the reference keeps it (the high byte carries over harmlessly: the slot
    // function is evaluated on the full 32-bit value there, so it is here)
    unsigned win = n >= 2 ? ((unsigned)in[0] << 8) | in[1] : 0;
    while (i + 2 < n) {
        win = (win << 8) | in[i + 2];
        const unsigned s = lzf_slot(win);
        const size_t cand = last[s];
        last[s] = (unsigned)i;
        const size_t dist_m1 = i - cand - 1;
        if (cand > 0 && cand < i && dist_m1 < 8192 && in[cand] == in[i] && in[cand + 1] == in[i + 1] && in[cand + 2] == in[i + 2]) {
            if (!w.room_for_match()) return 0;
            size_t limit = n - i - 2; // the match may not use the last two bytes' successors
            if (limit > 264) limit = 264;
            // As the reference: when more than 16 bytes remain, 16 are compared without looking at the limit
            // (so a match that ends within two bytes of it may overshoot it), then the bounded loop.
            size_t len = 2;
            bool differs = false;
            if (limit > 16)
                for (int c = 0; c < 16 && !differs; c++) {
                    len++;
                    differs = in[cand + len] != in[i + len];
                }
            if (!differs) {
                do
                    len++;
                while (len < limit && in[cand + len] == in[i + len]);
            }
            w.match(dist_m1, len - 2);
            i += len;
            if (i + 2 >= n) break;
            // only the last position of the match enters the table
            --i;
            win = lzf_h3(in + i);
            last[lzf_slot(win)] = (unsigned)i;
            ++i;
        } else {
            w.literal(in[i++]);
            if (!w.ok) return 0;
        }
    }
    if (w.at + 3 > cap) return 0;
    while (i < n) {
        w.literal(in[i++]);
        if (!w.ok) return 0;
    }
    w.close_run();
    return (unsigned)w.at;
}

unsigned lzf_decode(const unsigned char* in, size_t n, unsigned char* out, size_t cap)
{
    size_t i = 0, o = 0;
    while (i < n) {
        const unsigned c = in[i++];
        if (c < 32) {
            const size_t run = c + 1;
            if (o + run > cap || i + run > n) return 0;
            memcpy(out + o, in + i, run);
            o += run, i += run;
            continue;
        }
        size_t len = c >> 5;
        if (i >= n) return 0;
        if (len == 7) {
            len += in[i++];
            if (i >= n) return 0;
        }
        const size_t dist = ((size_t)(c & 31) << 8) + in[i++] + 1;
        len += 2;
        if (o + len > cap || dist > o) return 0;
        for (size_t k = 0, from = o - dist; k < len; k++) out[o++] = out[from++]; // overlap intended
    }
    return (unsigned)o;
}

} // namespace

// ---- PCD v0.7 ------------------------------------------------------------------------------------------------
namespace {

int io_fail(int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    pcp_set_global_error(buf);
    return code;
}

std::vector<std::string> split_ws(const std::string& line)
{
    std::vector<std::string> t;
    std::istringstream ss(line);
    std::string w;
    while (ss >> w) t.push_back(w);
    return t;
}

// header lines until DATA (pcd_helper.cpp:120-300); returns the offset of the payload
int parse_header(FILE* f, pcp_pcd_header* h)
{
    memset(h, 0, sizeof *h);
    h->viewpoint[3] = 1.0;
    bool have_points = false, have_data = false;
    std::string line;
    while (!have_data) {
        line.clear();
        int c;
        while ((c = fgetc(f)) != EOF && c != '\n') line.push_back((char)c);
        if (c == EOF && line.empty()) break;
        const std::vector<std::string> t = split_ws(line);
        if (t.empty() || t[0][0] == '#') continue;
        const std::string& key = t[0];
        const size_t nv = t.size() - 1;
        if (key == "VERSION") continue;
        if (key == "FIELDS" || key == "COLUMNS") {
            if (nv > PCP_PCD_MAX_FIELDS) return io_fail(PCP_ERR_RANGE, "PCD header: too many FIELDS");
            h->nfields = (int32_t)nv;
            for (size_t i = 0; i < nv; i++) {
                snprintf(h->names[i], sizeof h->names[i], "%s", t[i + 1].c_str());
                h->size[i] = 4, h->type[i] = 'F', h->count[i] = 1; // defaults of the reader
            }
        } else if (key == "SIZE") {
            if ((int32_t)nv != h->nfields) return io_fail(PCP_ERR_INVALID, "PCD header: SIZE and FIELDS differ in length");
            for (size_t i = 0; i < nv; i++) h->size[i] = atoi(t[i + 1].c_str());
        } else if (key == "TYPE") {
            if ((int32_t)nv != h->nfields) return io_fail(PCP_ERR_INVALID, "PCD header: TYPE and FIELDS differ in length");
            for (size_t i = 0; i < nv; i++) h->type[i] = t[i + 1][0];
        } else if (key == "COUNT") {
            if ((int32_t)nv != h->nfields) return io_fail(PCP_ERR_INVALID, "PCD header: COUNT and FIELDS differ in length");
            for (size_t i = 0; i < nv; i++) h->count[i] = atoi(t[i + 1].c_str());
        } else if (key == "WIDTH") {
            if (nv) h->width = (uint32_t)strtoul(t[1].c_str(), nullptr, 10);
        } else if (key == "HEIGHT") {
            if (nv) h->height = (uint32_t)strtoul(t[1].c_str(), nullptr, 10);
        } else if (key == "VIEWPOINT") {
            if (nv < 7) return io_fail(PCP_ERR_INVALID, "PCD header: VIEWPOINT needs 7 values (tx ty tz qw qx qy qz)");
            for (int i = 0; i < 7; i++) h->viewpoint[i] = strtod(t[i + 1].c_str(), nullptr);
        } else if (key == "POINTS") {
            if (nv) h->points = strtoull(t[1].c_str(), nullptr, 10), have_points = true;
        } else if (key == "DATA") {
            if (!nv) return io_fail(PCP_ERR_INVALID, "PCD header: DATA without a type");
            if (t[1].compare(0, 17, "binary_compressed") == 0) h->data_type = PCP_PCD_BINARY_COMPRESSED;
            else if (t[1].compare(0, 6, "binary") == 0) h->data_type = PCP_PCD_BINARY;
            else if (t[1].compare(0, 5, "ascii") == 0) h->data_type = PCP_PCD_ASCII;
            else return io_fail(PCP_ERR_INVALID, "PCD header: unknown DATA type '%s'", t[1].c_str());
            h->data_offset = (uint64_t)ftell(f);
            have_data = true;
        }
    }
    if (!have_data) return io_fail(PCP_ERR_INVALID, "PCD header: no DATA line");
    // unorganised defaults and the consistency check of pcd_helper.cpp:300-330
    if (!have_points) h->points = (uint64_t)h->width * (h->height ? h->height : 1);
    if (h->width == 0 && h->height == 0) h->width = (uint32_t)h->points, h->height = 1;
    if (h->height == 0) {
        h->height = 1;
        if (h->width == 0) h->width = (uint32_t)h->points;
    } else if (h->width == 0 && h->points != 0) {
        return io_fail(PCP_ERR_INVALID, "PCD header: HEIGHT given but no WIDTH");
    }
    if ((uint64_t)h->width * h->height != h->points) return io_fail(PCP_ERR_INVALID, "PCD header: HEIGHT x WIDTH != number of points");
    // A header is untrusted input: every size below feeds an allocation or a device-side stride.  Field sizes the
    // format knows (1, 2, 4, 8), positive bounded counts, the three type letters, floats of 4 or 8 bytes; and a binary
    // payload must fit into the file (a compressed one is checked against its own size words by the readers).
    uint64_t fsize = 0;
    for (int i = 0; i < h->nfields; i++) {
        const int sz = h->size[i], ct = h->count[i];
        if (sz != 1 && sz != 2 && sz != 4 && sz != 8) return io_fail(PCP_ERR_INVALID, "PCD header: SIZE %d of field '%s' is not 1, 2, 4 or 8", sz, h->names[i]);
        if (ct < 1 || ct > (1 << 20)) return io_fail(PCP_ERR_INVALID, "PCD header: COUNT %d of field '%s' out of range", ct, h->names[i]);
        if (h->type[i] != 'I' && h->type[i] != 'U' && h->type[i] != 'F') return io_fail(PCP_ERR_INVALID, "PCD header: TYPE '%c' of field '%s' is not I, U or F", h->type[i], h->names[i]);
        if (h->type[i] == 'F' && sz != 4 && sz != 8) return io_fail(PCP_ERR_INVALID, "PCD header: float field '%s' of %d bytes", h->names[i], sz);
        fsize += (uint64_t)sz * (uint64_t)ct;
    }
    if (h->points > (1ull << 40) || fsize > (1ull << 24)) return io_fail(PCP_ERR_RANGE, "PCD header: %llu points of %llu bytes are beyond this reader", (unsigned long long)h->points, (unsigned long long)fsize);
    if (h->data_type == PCP_PCD_BINARY) {
        const long here = ftell(f);
        fseek(f, 0, SEEK_END);
        const uint64_t file_size = (uint64_t)ftell(f);
        fseek(f, here, SEEK_SET);
        if (h->data_offset + h->points * fsize > file_size)
            return io_fail(PCP_ERR_INVALID, "PCD header: %llu points of %llu bytes do not fit into the file (%llu bytes after the header)",
                           (unsigned long long)h->points, (unsigned long long)fsize, (unsigned long long)(file_size - h->data_offset));
    }
    return PCP_OK;
}

// extern "C" entry points must not let an exception out (a failed allocation inside a std::vector, a worker thread)
#define PCP_IO_GUARD_BEGIN try {
#define PCP_IO_GUARD_END                                                                                              \
    }                                                                                                                 \
    catch (const std::bad_alloc&) { return io_fail(PCP_ERR_NOMEM, "PCD: out of host memory"); }                     \
    catch (const std::exception& e) { return io_fail(PCP_ERR_INVALID, "PCD: %s", e.what()); }

// the members a record can receive: name, PCD type, element size, offset in a PointXYZRGBA-shaped record
struct Member {
    const char* name;
    char type;
    int size, offset;
};
const Member kMembers[5] = {{"x", 'F', 8, 0}, {"y", 'F', 8, 8}, {"z", 'F', 8, 16}, {"rgba", 'U', 4, 32}, {"stamp_id", 'U', 4, 36}};
int member_count(int point_kind) { return point_kind == PCP_PCD_POINT_XYZ ? 3 : 5; }

// field of the file -> member of the record, matched on name + type + size + count (point_traits.h:271-279); -1: ignored
void match_fields(const pcp_pcd_header& h, int point_kind, int map[PCP_PCD_MAX_FIELDS])
{
    for (int i = 0; i < h.nfields; i++) {
        map[i] = -1;
        for (int m = 0; m < member_count(point_kind); m++)
            if (!strcmp(h.names[i], kMembers[m].name) && h.type[i] == kMembers[m].type && h.size[i] == kMembers[m].size && h.count[i] == 1)
                map[i] = m;
    }
}

void init_records(char* pts, size_t n, size_t stride, int point_kind)
{
    for (size_t i = 0; i < n; i++) {
        char* r = pts + i * stride;
        memset(r, 0, stride < 48 ? stride : 48); // PointXYZRGBA(): x = y = z = 0, rgba = 0 (point_type.h:86-91)
        if (stride >= 32) {
            const double one = 1.0;
            memcpy(r + 24, &one, 8); // data[3] = 1
        }
        (void)point_kind;
    }
}

bool finite_value(const char* p, char type, int size)
{
    if (type != 'F') return true;
    if (size == 8) {
        double v;
        memcpy(&v, p, 8);
        return std::isfinite(v);
    }
    float v;
    memcpy(&v, p, 4);
    return std::isfinite(v);
}

// copyStringValue (pcd_helper.h): "nan" -> quiet NaN, otherwise the stream conversion to the field's type
void store_ascii(const std::string& tok, char type, int size, char* dst)
{
    if (type == 'F') {
        const double v = (tok == "nan") ? std::nan("") : strtod(tok.c_str(), nullptr);
        if (size == 8) memcpy(dst, &v, 8);
        else {
            const float f = (float)v;
            memcpy(dst, &f, 4);
        }
    } else if (type == 'U') {
        const unsigned long long v = strtoull(tok.c_str(), nullptr, 10);
        memcpy(dst, &v, size); // little endian
    } else {
        const long long v = strtoll(tok.c_str(), nullptr, 10);
        memcpy(dst, &v, size);
    }
}

std::string make_header(int point_kind, uint32_t width, uint32_t height, size_t n, const char* data)
{
    // generateHeader<PointT> (pcd_helper.h:323-371): registered fields only, no padding entries
    std::ostringstream o;
    o.imbue(std::locale::classic());
    o << "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS";
    const int nm = member_count(point_kind);
    for (int m = 0; m < nm; m++) o << " " << kMembers[m].name;
    o << "\nSIZE";
    for (int m = 0; m < nm; m++) o << " " << kMembers[m].size;
    o << "\nTYPE";
    for (int m = 0; m < nm; m++) o << " " << kMembers[m].type;
    o << "\nCOUNT";
    for (int m = 0; m < nm; m++) o << " 1";
    o << "\nWIDTH " << width << "\nHEIGHT " << height << "\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS " << n << "\nDATA " << data << "\n";
    return o.str();
}

} // namespace

extern "C" {

unsigned int pcp_lzf_compress(const void* in, unsigned int in_len, void* out, unsigned int out_len)
{
    if (!in || !out) return 0;
    return lzf_encode(static_cast<const unsigned char*>(in), in_len, static_cast<unsigned char*>(out), out_len);
}

unsigned int pcp_lzf_decompress(const void* in, unsigned int in_len, void* out, unsigned int out_len)
{
    if (!in || !out) return 0;
    return lzf_decode(static_cast<const unsigned char*>(in), in_len, static_cast<unsigned char*>(out), out_len);
}

int pcp_pcd_read_header(const char* path, pcp_pcd_header* header)
{
    if (!path || !header) return io_fail(PCP_ERR_INVALID, "pcp_pcd_read_header: bad argument");
    FILE* f = fopen(path, "rb");
    if (!f) return io_fail(PCP_ERR_IO, "Could not open file %s", path);
    const int st = parse_header(f, header);
    fclose(f);
    return st;
}

static int pcd_load_impl(const char* path, int point_kind, void* pts, size_t cap, size_t stride, size_t* n_out, int* is_dense);
int pcp_pcd_load(const char* path, int point_kind, void* pts, size_t cap, size_t stride, size_t* n_out, int* is_dense)
{
    PCP_IO_GUARD_BEGIN
    return pcd_load_impl(path, point_kind, pts, cap, stride, n_out, is_dense);
    PCP_IO_GUARD_END
}
static int pcd_load_impl(const char* path, int point_kind, void* pts, size_t cap, size_t stride, size_t* n_out, int* is_dense)
{
    if (!path || !n_out || (cap && !pts) || stride < (size_t)(point_kind == PCP_PCD_POINT_XYZ ? 24 : 40))
        return io_fail(PCP_ERR_INVALID, "pcp_pcd_load: bad argument");
    *n_out = 0;
    if (is_dense) *is_dense = 1;
    FILE* f = fopen(path, "rb");
    if (!f) return io_fail(PCP_ERR_IO, "Could not open file %s", path);
    pcp_pcd_header h;
    int st = parse_header(f, &h);
    if (st != PCP_OK) {
        fclose(f);
        return st;
    }
    const size_t n = (size_t)h.points;
    *n_out = n;
    if (n > cap) {
        fclose(f);
        return io_fail(PCP_ERR_RANGE, "pcp_pcd_load: the file holds more points than the buffer");
    }
    int map[PCP_PCD_MAX_FIELDS];
    match_fields(h, point_kind, map);
    char* out = static_cast<char*>(pts);
    init_records(out, n, stride, point_kind);
    size_t fsize = 0; // bytes of one point in the file (padding fields "_" included for binary, excluded for compressed)
    for (int i = 0; i < h.nfields; i++) fsize += (size_t)h.size[i] * h.count[i];
    bool dense = true;
    fseek(f, (long)h.data_offset, SEEK_SET);
    if (h.data_type == PCP_PCD_ASCII) {
        // one point per non-empty line, tokens in field order (pcd_helper.cpp:880-960); "_" fields consume tokens
        std::string line;
        size_t idx = 0;
        while (idx < n) {
            line.clear();
            int c;
            while ((c = fgetc(f)) != EOF && c != '\n') line.push_back((char)c);
            if (c == EOF && line.empty()) break;
            const std::vector<std::string> t = split_ws(line);
            if (t.empty()) continue;
            size_t at = 0;
            for (int i = 0; i < h.nfields; i++) {
                if (map[i] >= 0) {
                    if (at >= t.size()) {
                        fclose(f);
                        return io_fail(PCP_ERR_INVALID, "PCD ascii payload: a line has fewer values than fields");
                    }
                    store_ascii(t[at], h.type[i], h.size[i], out + idx * stride + kMembers[map[i]].offset);
                }
                at += (size_t)h.count[i];
            }
            idx++;
        }
        fclose(f);
        if (idx != n) return io_fail(PCP_ERR_INVALID, "Number of points read is different than expected");
        return PCP_OK; // is_dense stays true for ascii, as in the reference (pcd_helper.cpp:1118-1120)
    }
    std::vector<unsigned char> raw;
    if (h.data_type == PCP_PCD_BINARY) {
        raw.resize(n * fsize);
        if (n && fread(raw.data(), 1, raw.size(), f) != raw.size()) {
            fclose(f);
            return io_fail(PCP_ERR_IO, "PCD binary payload: file shorter than its header says");
        }
        fclose(f);
        for (size_t p = 0; p < n; p++) {
            size_t off = 0;
            for (int i = 0; i < h.nfields; i++) {
                const size_t bytes = (size_t)h.size[i] * h.count[i];
                const unsigned char* src = raw.data() + p * fsize + off;
                if (map[i] >= 0) memcpy(out + p * stride + kMembers[map[i]].offset, src, bytes);
                if (strcmp(h.names[i], "_")) // padding is not looked at
                    for (int c = 0; c < h.count[i]; c++) dense &= finite_value((const char*)src + (size_t)c * h.size[i], h.type[i], h.size[i]);
                off += bytes;
            }
        }
    } else {
        // u32 compressed size, u32 uncompressed size, one LZF stream of the fields as planes: all x, all y, ...
        // ("_" fields are not stored), pcd_helper.cpp:1010-1090
        unsigned int sizes[2];
        if (fread(sizes, 4, 2, f) != 2) {
            fclose(f);
            return io_fail(PCP_ERR_IO, "PCD binary_compressed payload: missing size words");
        }
        std::vector<unsigned char> comp(sizes[0]);
        if (sizes[0] && fread(comp.data(), 1, comp.size(), f) != comp.size()) {
            fclose(f);
            return io_fail(PCP_ERR_IO, "PCD binary_compressed payload: file shorter than its header says");
        }
        fclose(f);
        raw.resize(sizes[1]);
        if (lzf_decode(comp.data(), comp.size(), raw.data(), raw.size()) != sizes[1])
            return io_fail(PCP_ERR_INVALID, "Size of decompressed lzf data does not match value stored in PCD header");
        size_t plane = 0;
        for (int i = 0; i < h.nfields; i++) {
            if (!strcmp(h.names[i], "_")) continue;
            const size_t bytes = (size_t)h.size[i] * h.count[i];
            if (plane + bytes * n > raw.size()) return io_fail(PCP_ERR_INVALID, "PCD binary_compressed payload: planes exceed the uncompressed size");
            for (size_t p = 0; p < n; p++) {
                const unsigned char* src = raw.data() + plane + p * bytes;
                if (map[i] >= 0) memcpy(out + p * stride + kMembers[map[i]].offset, src, bytes);
                for (int c = 0; c < h.count[i]; c++) dense &= finite_value((const char*)src + (size_t)c * h.size[i], h.type[i], h.size[i]);
            }
            plane += bytes * n;
        }
    }
    if (is_dense) *is_dense = dense ? 1 : 0;
    return PCP_OK;
}

static int pcd_save_impl(const char* path, int point_kind, const void* pts, size_t n, size_t stride, uint32_t width, uint32_t height,
                 int mode, int precision)
;
int pcp_pcd_save(const char* path, int point_kind, const void* pts, size_t n, size_t stride, uint32_t width, uint32_t height,
                 int mode, int precision)
{
    PCP_IO_GUARD_BEGIN
    return pcd_save_impl(path, point_kind, pts, n, stride, width, height, mode, precision);
    PCP_IO_GUARD_END
}
static int pcd_save_impl(const char* path, int point_kind, const void* pts, size_t n, size_t stride, uint32_t width, uint32_t height,
                 int mode, int precision)
{
    if (!path || (n && !pts) || stride < (size_t)(point_kind == PCP_PCD_POINT_XYZ ? 24 : 40) || mode < 0 || mode > 2)
        return io_fail(PCP_ERR_INVALID, "pcp_pcd_save: bad argument");
    if (n == 0) return io_fail(PCP_ERR_INVALID, "Input point cloud has no data!"); // every writer refuses an empty cloud
    if (width == 0 && height == 0) width = (uint32_t)n, height = 1;
    if ((size_t)width * height != n) return io_fail(PCP_ERR_INVALID, "Number of points different than width * height!");
    const int nm = member_count(point_kind);
    size_t fsize = 0;
    for (int m = 0; m < nm; m++) fsize += kMembers[m].size;
    const char* in = static_cast<const char*>(pts);
    FILE* f = fopen(path, "wb");
    if (!f) return io_fail(PCP_ERR_IO, "Could not open file %s for writing", path);
    int st = PCP_OK;
    if (mode == PCP_PCD_ASCII) {
        // writeASCII (pcd_helper.h:947-1090): stream formatting with `precision` significant digits, "nan" for NaN
        const std::string hd = make_header(point_kind, width, height, n, "ascii");
        fwrite(hd.data(), 1, hd.size(), f);
        std::ostringstream s;
        s.imbue(std::locale::classic());
        s.precision(precision > 0 ? precision : 8);
        for (size_t p = 0; p < n; p++) {
            s.str("");
            for (int m = 0; m < nm; m++) {
                const char* src = in + p * stride + kMembers[m].offset;
                if (kMembers[m].type == 'F') {
                    double v;
                    memcpy(&v, src, 8);
                    if (std::isnan(v)) s << "nan";
                    else s << v;
                } else {
                    uint32_t v;
                    memcpy(&v, src, 4);
                    s << v;
                }
                if (m + 1 < nm) s << " ";
            }
            s << "\n";
            const std::string l = s.str();
            fwrite(l.data(), 1, l.size(), f);
        }
    } else {
        // the registered fields only, packed (writeBinary, pcd_helper.h:518-580), or as planes through LZF
        // (writeBinaryCompressed, pcd_helper.h:660-735)
        std::vector<unsigned char> packed(n * fsize);
        if (mode == PCP_PCD_BINARY) {
            unsigned char* o = packed.data();
            for (size_t p = 0; p < n; p++)
                for (int m = 0; m < nm; m++) {
                    memcpy(o, in + p * stride + kMembers[m].offset, kMembers[m].size);
                    o += kMembers[m].size;
                }
            const std::string hd = make_header(point_kind, width, height, n, "binary");
            fwrite(hd.data(), 1, hd.size(), f);
            fwrite(packed.data(), 1, packed.size(), f);
        } else {
            if (packed.size() > 0xffffffffull) st = io_fail(PCP_ERR_RANGE, "binary_compressed holds at most 4 GiB of fields (32-bit size words)");
            size_t plane = 0;
            for (int m = 0; m < nm && st == PCP_OK; m++) {
                for (size_t p = 0; p < n; p++) memcpy(packed.data() + plane + p * kMembers[m].size, in + p * stride + kMembers[m].offset, kMembers[m].size);
                plane += n * kMembers[m].size;
            }
            if (st == PCP_OK) {
                const unsigned int room = (unsigned int)((float)packed.size() * 1.5f); // the reference's buffer
                std::vector<unsigned char> comp((size_t)room + 8);
                const unsigned int csize = lzf_encode(packed.data(), packed.size(), comp.data() + 8, room);
                if (!csize) st = io_fail(PCP_ERR_INVALID, "Error during compression!");
                else {
                    const unsigned int usize = (unsigned int)packed.size();
                    memcpy(comp.data(), &csize, 4);
                    memcpy(comp.data() + 4, &usize, 4);
                    const std::string hd = make_header(point_kind, width, height, n, "binary_compressed");
                    fwrite(hd.data(), 1, hd.size(), f);
                    fwrite(comp.data(), 1, (size_t)csize + 8, f);
                }
            }
        }
    }
    if (ferror(f) && st == PCP_OK) st = io_fail(PCP_ERR_IO, "write error on %s", path);
    fclose(f);
    return st;
}

} // extern "C"

// ---- pipelined loader ----------------------------------------------------------------------------------------------
namespace {

// where the values of one file field live in the uploaded payload: value c of point p at
// raw + off + p * pstride + c * size (packed records: pstride = bytes of a point; planes: pstride = size * count)
struct PcdDevField {
    unsigned long long off;
    unsigned int pstride;
    int size, count;
    int dst;      // byte offset of the matched member in the record, -1: not stored
    int is_float; // checked for is_dense
};
struct PcdDevLayout {
    int nfields;
    PcdDevField f[PCP_PCD_MAX_FIELDS];
};

// little-endian value of `size` <= 8 bytes; one load when the address allows it (the layouts the writers produce)
__device__ __forceinline__ unsigned long long pcd_load_le(const unsigned char* src, int size)
{
    if (size == 8 && (reinterpret_cast<size_t>(src) & 7) == 0) return *reinterpret_cast<const unsigned long long*>(src);
    if (size == 4 && (reinterpret_cast<size_t>(src) & 3) == 0) return *reinterpret_cast<const unsigned int*>(src);
    unsigned long long v = 0;
    for (int b = 0; b < size && b < 8; b++) v |= (unsigned long long)src[b] << (8 * b);
    return v;
}

__global__ void __launch_bounds__(256) pcd_assemble_kernel(const unsigned char* __restrict__ raw, size_t n, PcdDevLayout L,
                                                           char* __restrict__ out, size_t stride, int* __restrict__ not_dense)
{
    const size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (p >= n) return;
    char* r = out + p * stride;
    // PointXYZRGBA(): x = y = z = 0, data[3] = 1, rgba = 0 (point_type.h:86-91)
    unsigned long long rec[6] = {0ull, 0ull, 0ull, 0x3ff0000000000000ull, 0ull, 0ull};
    const int words = (int)(stride < 48 ? stride : 48) / 8;
    bool dense = true;
    for (int i = 0; i < L.nfields; i++) {
        const PcdDevField f = L.f[i];
        const unsigned char* src = raw + f.off + p * (size_t)f.pstride;
        if (f.dst >= 0) { // count == 1; members are 4 or 8 bytes at aligned offsets of the record
            const unsigned long long v = pcd_load_le(src, f.size);
            if (f.size == 8) rec[f.dst >> 3] = v;
            else if (f.dst & 4) rec[f.dst >> 3] = (rec[f.dst >> 3] & 0xffffffffull) | (v << 32);
            else rec[f.dst >> 3] = (rec[f.dst >> 3] & ~0xffffffffull) | v;
        }
        if (f.is_float)
            for (int c = 0; c < f.count; c++) {
                const unsigned long long v = pcd_load_le(src + (size_t)c * f.size, f.size);
                if (f.size == 8) dense &= isfinite(__longlong_as_double((long long)v));
                else if (f.size == 4) dense &= isfinite(__uint_as_float((unsigned)v));
            }
    }
    for (int w = 0; w < words; w++) reinterpret_cast<unsigned long long*>(r)[w] = rec[w];
    if (!dense) *not_dense = 1;
}

struct PcdLoaded {
    int status = PCP_OK;
    std::string error;
    pcp_pcd_header h;
    std::vector<unsigned char> raw;   // payload as uploaded (binary: packed records; compressed: planes)
    std::vector<unsigned char> recs;  // ascii: records parsed on the host
    PcdDevLayout L;
};

// host half of one file: payload into memory, decoded if compressed, layout of its fields
void load_one(const char* path, int point_kind, size_t stride, PcdLoaded& out)
{
    auto fail = [&](int code, const char* msg) {
        out.status = code;
        out.error = std::string(msg) + " (" + path + ")";
    };
    FILE* f = fopen(path, "rb");
    if (!f) return fail(PCP_ERR_IO, "Could not open file");
    pcp_pcd_header& h = out.h;
    if (parse_header(f, &h) != PCP_OK) {
        fclose(f);
        return fail(PCP_ERR_INVALID, "bad PCD header");
    }
    const size_t n = (size_t)h.points;
    int map[PCP_PCD_MAX_FIELDS];
    match_fields(h, point_kind, map);
    size_t fsize = 0;
    for (int i = 0; i < h.nfields; i++) fsize += (size_t)h.size[i] * h.count[i];
    fseek(f, (long)h.data_offset, SEEK_SET);
    if (h.data_type == PCP_PCD_ASCII) { // parsed on the host by the single-file reader
        fclose(f);
        out.recs.resize(n * stride);
        size_t got = 0;
        int dense = 1;
        const int st = pcp_pcd_load(path, point_kind, out.recs.data(), n, stride, &got, &dense);
        if (st != PCP_OK) return fail(st, "ascii payload");
        return;
    }
    PcdDevLayout& L = out.L;
    L.nfields = 0;
    if (h.data_type == PCP_PCD_BINARY) {
        out.raw.resize(n * fsize);
        const bool ok = !n || fread(out.raw.data(), 1, out.raw.size(), f) == out.raw.size();
        fclose(f);
        if (!ok) return fail(PCP_ERR_IO, "PCD binary payload: file shorter than its header says");
        size_t off = 0;
        for (int i = 0; i < h.nfields; i++) {
            const bool pad = !strcmp(h.names[i], "_");
            if (!pad) {
                PcdDevField& d = L.f[L.nfields++];
                d.off = off, d.pstride = (unsigned)fsize, d.size = h.size[i], d.count = h.count[i];
                d.dst = map[i] >= 0 ? kMembers[map[i]].offset : -1;
                d.is_float = h.type[i] == 'F';
            }
            off += (size_t)h.size[i] * h.count[i];
        }
        return;
    }
    unsigned int sizes[2];
    if (fread(sizes, 4, 2, f) != 2) {
        fclose(f);
        return fail(PCP_ERR_IO, "PCD binary_compressed payload: missing size words");
    }
    std::vector<unsigned char> comp(sizes[0]);
    const bool ok = !sizes[0] || fread(comp.data(), 1, comp.size(), f) == comp.size();
    fclose(f);
    if (!ok) return fail(PCP_ERR_IO, "PCD binary_compressed payload: file shorter than its header says");
    out.raw.resize(sizes[1]);
    if (lzf_decode(comp.data(), comp.size(), out.raw.data(), out.raw.size()) != sizes[1])
        return fail(PCP_ERR_INVALID, "Size of decompressed lzf data does not match value stored in PCD header");
    size_t plane = 0;
    for (int i = 0; i < h.nfields; i++) {
        if (!strcmp(h.names[i], "_")) continue;
        const size_t bytes = (size_t)h.size[i] * h.count[i];
        if (plane + bytes * n > out.raw.size()) return fail(PCP_ERR_INVALID, "PCD binary_compressed payload: planes exceed the uncompressed size");
        PcdDevField& d = L.f[L.nfields++];
        d.off = plane, d.pstride = (unsigned)bytes, d.size = h.size[i], d.count = h.count[i];
        d.dst = map[i] >= 0 ? kMembers[map[i]].offset : -1;
        d.is_float = h.type[i] == 'F';
        plane += bytes * n;
    }
}

} // namespace

extern "C" int pcp_pcd_load_many(pcp_ctx* ctx, const char* const* paths, size_t nfiles, int point_kind, void* pts, size_t cap,
                                 size_t stride, int threads, size_t* offsets, int* is_dense)
{
    if (!ctx || !offsets || (nfiles && !paths) || (cap && !pts) || stride < (size_t)(point_kind == PCP_PCD_POINT_XYZ ? 24 : 40) ||
        stride % 8)
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_pcd_load_many: bad argument");
    if (is_dense) *is_dense = 1;
    offsets[0] = 0;
    // sizes first: the offsets of every file's records are known before any payload is read
    for (size_t i = 0; i < nfiles; i++) {
        pcp_pcd_header h;
        if (pcp_pcd_read_header(paths[i], &h) != PCP_OK) return pcp_fail(ctx, PCP_ERR_IO, "%s", pcp_last_error(nullptr));
        offsets[i + 1] = offsets[i] + (size_t)h.points;
    }
    const size_t total = offsets[nfiles];
    if (total > cap) return pcp_fail(ctx, PCP_ERR_RANGE, "pcp_pcd_load_many: the files hold %zu points, the buffer %zu", total, cap);
    if (total == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    OutStage out;
    PCP_TRY(out.open(ctx, pts, total * stride));
    DevBuf<int> nd;
    PCP_TRY(nd.alloc(ctx, 1));
    PCP_CUDA_TRY(ctx, cudaMemsetAsync(nd.p, 0, sizeof(int), ctx->stream));
    // host threads take files in order; the calling thread uploads and assembles them in order as they complete
    if (threads < 1) threads = (int)std::thread::hardware_concurrency();
    if (threads < 1) threads = 1;
    if ((size_t)threads > nfiles) threads = (int)nfiles;
    std::vector<std::unique_ptr<PcdLoaded>> loaded(nfiles);
    std::vector<std::promise<void>> ready(nfiles);
    std::atomic<size_t> next(0);
    std::atomic<bool> stop(false);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++)
        pool.emplace_back([&]() {
            for (;;) {
                const size_t i = next.fetch_add(1);
                if (i >= nfiles) return;
                loaded[i].reset(new PcdLoaded);
                if (!stop.load()) {
                    try {
                        load_one(paths[i], point_kind, stride, *loaded[i]);
                    } catch (const std::bad_alloc&) {
                        loaded[i]->status = PCP_ERR_NOMEM, loaded[i]->error = std::string("out of host memory (") + paths[i] + ")";
                    } catch (const std::exception& e) {
                        loaded[i]->status = PCP_ERR_INVALID, loaded[i]->error = std::string(e.what()) + " (" + paths[i] + ")";
                    }
                }
                ready[i].set_value();
            }
        });
    int status = PCP_OK;
    std::string err;
    for (size_t i = 0; i < nfiles; i++) {
        ready[i].get_future().wait();
        std::unique_ptr<PcdLoaded> L = std::move(loaded[i]);
        if (status != PCP_OK) continue;
        if (L->status != PCP_OK) {
            status = L->status, err = L->error;
            stop.store(true);
            continue;
        }
        const size_t n = (size_t)L->h.points;
        if (n != offsets[i + 1] - offsets[i]) {
            status = PCP_ERR_IO, err = std::string("file changed while loading: ") + paths[i];
            stop.store(true);
            continue;
        }
        if (!n) continue;
        char* dst = static_cast<char*>(out.dptr) + offsets[i] * stride;
        cudaError_t e = cudaSuccess;
        if (!L->recs.empty()) {
            e = cudaMemcpyAsync(dst, L->recs.data(), n * stride, cudaMemcpyHostToDevice, ctx->stream);
        } else {
            void* draw = nullptr;
            if (pcp_arena_alloc(ctx, L->raw.size(), &draw) != PCP_OK) {
                status = PCP_ERR_CUDA, err = "device staging allocation failed";
                stop.store(true);
                continue;
            }
            e = cudaMemcpyAsync(draw, L->raw.data(), L->raw.size(), cudaMemcpyHostToDevice, ctx->stream);
            if (e == cudaSuccess) {
                pcd_assemble_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(static_cast<const unsigned char*>(draw), n, L->L,
                                                                                     dst, stride, nd.p);
                ctx->launches++;
                e = cudaGetLastError();
            }
            pcp_arena_free(ctx, draw); // stream-ordered: after the kernel
        }
        if (e != cudaSuccess) {
            status = PCP_ERR_CUDA, err = cudaGetErrorString(e);
            stop.store(true);
        }
    }
    for (std::thread& t : pool) t.join();
    if (status != PCP_OK) {
        cudaStreamSynchronize(ctx->stream);
        return pcp_fail(ctx, status, "pcp_pcd_load_many: %s", err.c_str());
    }
    int not_dense = 0;
    PCP_TRY(pcp_readback(ctx, &not_dense, nd.p, sizeof(int)));
    if (is_dense) *is_dense = not_dense ? 0 : 1;
    return out.commit(total * stride);
}
