// 8-bit RGBA, non-interlaced PNG encoder on zlib (replaces the libpng calls of
// gpu/rt.cpp:14-52; libpng is absent from this image).  Only the decoded pixels matter
// for parity; filter 0 (None) on every row, one IDAT chunk.
// Rows are deflated in parallel slabs: each slab is an independent raw
// deflate stream ended with Z_SYNC_FLUSH / Z_FINISH, concatenated inside one zlib
// stream whose Adler-32 is combined from the per-slab checksums.
#include <zlib.h>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>
#include "rt_host.h"

namespace {

void put_be32(std::vector<unsigned char> &v, uint32_t x) {
  v.push_back((unsigned char)(x >> 24));
  v.push_back((unsigned char)(x >> 16));
  v.push_back((unsigned char)(x >> 8));
  v.push_back((unsigned char)x);
}

bool write_chunk(FILE *f, const char type[4], const unsigned char *data, size_t len) {
  unsigned char hdr[8] = {(unsigned char)(len >> 24), (unsigned char)(len >> 16), (unsigned char)(len >> 8),
                          (unsigned char)len, (unsigned char)type[0], (unsigned char)type[1],
                          (unsigned char)type[2], (unsigned char)type[3]};
  uLong crc = crc32(0L, hdr + 4, 4);
  if (len) crc = crc32(crc, data, (uInt)len);
  unsigned char tail[4] = {(unsigned char)(crc >> 24), (unsigned char)(crc >> 16), (unsigned char)(crc >> 8),
                           (unsigned char)crc};
  return fwrite(hdr, 1, 8, f) == 8 && (len == 0 || fwrite(data, 1, len, f) == len) && fwrite(tail, 1, 4, f) == 4;
}

struct Slab {
  int row0 = 0, row1 = 0;
  std::vector<unsigned char> out;
  uLong adler = 1;
  size_t raw_len = 0;
  bool ok = true;
};

void deflate_slab(Slab &s, const unsigned char *buffer, int width, ptrdiff_t stride, bool last) {
  const size_t row_bytes = (size_t)width * 4 + 1;
  std::vector<unsigned char> raw((size_t)(s.row1 - s.row0) * row_bytes);
  for (int r = s.row0; r < s.row1; ++r) {
    unsigned char *dst = raw.data() + (size_t)(r - s.row0) * row_bytes;
    dst[0] = 0;  // filter type None
    memcpy(dst + 1, buffer + (ptrdiff_t)r * stride, (size_t)width * 4);
  }
  s.raw_len = raw.size();
  s.adler = adler32(1L, raw.data(), (uInt)raw.size());
  z_stream z;
  memset(&z, 0, sizeof z);
  if (deflateInit2(&z, Z_DEFAULT_COMPRESSION, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY) != Z_OK) {
    s.ok = false;
    return;
  }
  s.out.resize(deflateBound(&z, (uLong)raw.size()) + 16);
  z.next_in = raw.data();
  z.avail_in = (uInt)raw.size();
  z.next_out = s.out.data();
  z.avail_out = (uInt)s.out.size();
  int rc = deflate(&z, last ? Z_FINISH : Z_FULL_FLUSH);
  s.ok = last ? rc == Z_STREAM_END : rc == Z_OK;
  s.out.resize(s.out.size() - z.avail_out);
  deflateEnd(&z);
}

}  // namespace

extern "C" int rtb_write_png(const unsigned char *buffer, int width, int height, ptrdiff_t stride,
                             const char *filename) {
  if (!buffer || width <= 0 || height <= 0 || !filename) return -1;
  unsigned hw = std::thread::hardware_concurrency();
  int nslab = (int)(hw ? hw : 1);
  if (nslab > 64) nslab = 64;
  const int min_rows = 16;
  if (nslab > (height + min_rows - 1) / min_rows) nslab = (height + min_rows - 1) / min_rows;
  if ((size_t)width * 4 * (size_t)height < (1u << 20)) nslab = 1;
  std::vector<Slab> slabs((size_t)nslab);
  for (int i = 0; i < nslab; ++i) {
    slabs[i].row0 = (int)((long long)height * i / nslab);
    slabs[i].row1 = (int)((long long)height * (i + 1) / nslab);
  }
  std::vector<std::thread> pool;
  for (int i = 1; i < nslab; ++i)
    pool.emplace_back(deflate_slab, std::ref(slabs[i]), buffer, width, stride, i == nslab - 1);
  deflate_slab(slabs[0], buffer, width, stride, nslab == 1);
  for (auto &t : pool) t.join();

  std::vector<unsigned char> idat;
  idat.push_back(0x78);  // zlib header: deflate, 32 K window
  idat.push_back(0x9C);  // default compression, FCHECK
  uLong adler = 1;
  for (int i = 0; i < nslab; ++i) {
    if (!slabs[i].ok) return -1;
    idat.insert(idat.end(), slabs[i].out.begin(), slabs[i].out.end());
    adler = i == 0 ? slabs[0].adler : adler32_combine(adler, slabs[i].adler, (z_off_t)slabs[i].raw_len);
  }
  put_be32(idat, (uint32_t)adler);

  FILE *f = fopen(filename, "wb");
  if (!f) return -1;
  static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
  bool ok = fwrite(sig, 1, 8, f) == 8;
  std::vector<unsigned char> ihdr;
  put_be32(ihdr, (uint32_t)width);
  put_be32(ihdr, (uint32_t)height);
  ihdr.push_back(8);  // bit depth
  ihdr.push_back(6);  // colour type RGBA
  ihdr.push_back(0);  // compression
  ihdr.push_back(0);  // filter method
  ihdr.push_back(0);  // no interlace
  ok = ok && write_chunk(f, "IHDR", ihdr.data(), ihdr.size());
  ok = ok && write_chunk(f, "IDAT", idat.data(), idat.size());
  ok = ok && write_chunk(f, "IEND", nullptr, 0);
  ok = (fclose(f) == 0) && ok;
  return ok ? 0 : -1;
}
