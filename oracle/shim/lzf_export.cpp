// C-linkage doorway onto the reference's own lzfCompress/lzfDecompress (lzf.h:62-88), so that tests can
// check the oracle's and the product's LZF against the real thing via ctypes.
#include "lzf.h"
extern "C" unsigned int ref_lzf_compress(const void* in, unsigned int in_len, void* out, unsigned int out_len)
{
    return cloud_blend_double::lzfCompress(in, in_len, out, out_len);
}
extern "C" unsigned int ref_lzf_decompress(const void* in, unsigned int in_len, void* out, unsigned int out_len)
{
    return cloud_blend_double::lzfDecompress(in, in_len, out, out_len);
}
