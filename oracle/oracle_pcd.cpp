// oracle_pcd.cpp -- TEST INFRASTRUCTURE (see oracle.h).
// LZF restated from the algorithm in /root/reference/lzf.cpp:86-415 (Marc Lehmann's liblzf, HLOG 13,
// the default non-VERY_FAST parse), index-based instead of pointer-based.  Pinned byte-for-byte against
// the reference's own lzf.cpp compiled into oracle/_ref/liblzf_ref.so (tests/test_oracle_lzf.py).
//
// stream format (lzf.cpp:75-82):
//   000LLLLL <L+1 bytes>          literal run of L+1 = 1..32 bytes
//   LLLooooo oooooooo             back reference, L+2 = 3..8 bytes, distance o+1
//   111ooooo LLLLLLLL oooooooo    back reference, L+9 bytes, distance o+1
#include "oracle.h"
#include <cstring>
#include <vector>

namespace {
const unsigned HLOG = 13;
const unsigned MAX_LIT = 1u << 5;
const unsigned MAX_OFF = 1u << 13;
const unsigned MAX_REF = (1u << 8) + (1u << 3);
inline unsigned slot_of(unsigned h) { return ((h >> (3 * 8 - HLOG)) - h) & ((1u << HLOG) - 1); }
} // namespace

extern "C" unsigned int orc_lzf_compress(const void* in_data, unsigned int in_len, void* out_data, unsigned int out_len)
{
    if (!in_len || !out_len) return 0;
    const unsigned char* in = static_cast<const unsigned char*>(in_data);
    unsigned char* out = static_cast<unsigned char*>(out_data);
    std::vector<unsigned> htab(1u << HLOG, 0u);
    long long n = in_len, cap = out_len;
    long long ip = 0, op = 1; // op=1: byte 0 is reserved for the first run header
    int lit = 0;
    unsigned hval = in_len >= 2 ? ((unsigned)in[0] << 8) | in[1] : 0;
    while (ip < n - 2) {
        hval = (hval << 8) | in[ip + 2];
        unsigned s = slot_of(hval);
        long long ref = htab[s];
        htab[s] = (unsigned)ip;
        long long off = ip - ref - 1;
        if (ref < ip && off < (long long)MAX_OFF && ref > 0 && in[ref + 2] == in[ip + 2] && in[ref] == in[ip] &&
            in[ref + 1] == in[ip + 1]) {
            long long len = 2;
            long long maxlen = n - ip - len;
            if (maxlen > (long long)MAX_REF) maxlen = MAX_REF;
            if (op + 3 + 1 >= cap)
                if (op - (lit ? 0 : 1) + 3 + 1 >= cap) return 0;
            out[op - lit - 1] = (unsigned char)(lit - 1); // close the literal run
            if (!lit) op--;                               // ...or drop its empty header
            bool stopped = false;
            if (maxlen > 16) {
                for (int c = 0; c < 16; c++) {
                    len++;
                    if (in[ref + len] != in[ip + len]) {
                        stopped = true;
                        break;
                    }
                }
            }
            if (!stopped) {
                do
                    len++;
                while (len < maxlen && in[ref + len] == in[ip + len]);
            }
            len -= 2; // now (#bytes matched) - 2
            ip++;
            if (len < 7) {
                out[op++] = (unsigned char)((off >> 8) + (len << 5));
            } else {
                out[op++] = (unsigned char)((off >> 8) + (7 << 5));
                out[op++] = (unsigned char)(len - 7);
            }
            out[op++] = (unsigned char)off;
            lit = 0;
            op++; // reserve next run header
            ip += len + 1;
            if (ip >= n - 2) break;
            --ip;
            hval = ((unsigned)in[ip] << 8) | in[ip + 1];
            hval = (hval << 8) | in[ip + 2];
            htab[slot_of(hval)] = (unsigned)ip;
            ip++;
        } else {
            if (op >= cap) return 0;
            lit++;
            out[op++] = in[ip++];
            if (lit == (int)MAX_LIT) {
                out[op - lit - 1] = (unsigned char)(lit - 1);
                lit = 0;
                op++;
            }
        }
    }
    if (op + 3 > cap) return 0;
    while (ip < n) {
        lit++;
        out[op++] = in[ip++];
        if (lit == (int)MAX_LIT) {
            out[op - lit - 1] = (unsigned char)(lit - 1);
            lit = 0;
            op++;
        }
    }
    out[op - lit - 1] = (unsigned char)(lit - 1);
    if (!lit) op--;
    return (unsigned int)op;
}

extern "C" unsigned int orc_lzf_decompress(const void* in_data, unsigned int in_len, void* out_data, unsigned int out_len)
{
    const unsigned char* in = static_cast<const unsigned char*>(in_data);
    unsigned char* out = static_cast<unsigned char*>(out_data);
    size_t ip = 0, op = 0, n = in_len, cap = out_len;
    while (ip < n) {
        unsigned ctrl = in[ip++];
        if (ctrl < (1u << 5)) { // literal run
            size_t run = ctrl + 1;
            if (op + run > cap) return 0; // E2BIG
            if (ip + run > n) return 0;   // EINVAL
            std::memcpy(out + op, in + ip, run);
            op += run;
            ip += run;
        } else { // back reference
            size_t len = ctrl >> 5;
            if (ip >= n) return 0;
            if (len == 7) {
                len += in[ip++];
                if (ip >= n) return 0;
            }
            size_t dist = ((size_t)(ctrl & 0x1f) << 8) + in[ip++] + 1;
            len += 2;
            if (op + len > cap) return 0;
            if (dist > op) return 0;
            size_t from = op - dist;
            for (size_t c = 0; c < len; c++) out[op++] = out[from++]; // overlapping copy is intended
        }
    }
    return (unsigned int)op;
}
