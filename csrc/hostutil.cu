// Host-side native helpers (no device code): CRC-32C for TensorBundle checkpoints, TFRecord and
// TensorBoard event framing.  The reference relies on TensorFlow's C++ runtime for these
// (tf.train.Saver, tf.summary.FileWriter, TFRecordDataset); SURVEY.md 5.4 / Appendix B.
#include <stddef.h>
#include <stdint.h>
#include <string.h>

namespace {

uint32_t g_table[8][256];
bool g_init = false;

void init_tables() {
  const uint32_t poly = 0x82F63B78u;  // Castagnoli, reflected
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1) ? (c >> 1) ^ poly : (c >> 1);
    g_table[0][i] = c;
  }
  for (uint32_t i = 0; i < 256; ++i) {
    uint32_t c = g_table[0][i];
    for (int t = 1; t < 8; ++t) {
      c = g_table[0][c & 0xff] ^ (c >> 8);
      g_table[t][i] = c;
    }
  }
  g_init = true;
}

}  // namespace

// crc = drn_crc32c(data, n, 0) ; chain with the previous return value to extend.
extern "C" uint32_t drn_crc32c(const void* data, size_t n, uint32_t crc) {
  if (!g_init) init_tables();
  const uint8_t* p = static_cast<const uint8_t*>(data);
  uint32_t c = crc ^ 0xffffffffu;
  while (n && (reinterpret_cast<uintptr_t>(p) & 7)) {
    c = g_table[0][(c ^ *p++) & 0xff] ^ (c >> 8);
    --n;
  }
  while (n >= 8) {
    uint64_t w;
    memcpy(&w, p, 8);
    w ^= c;
    c = g_table[7][w & 0xff] ^ g_table[6][(w >> 8) & 0xff] ^ g_table[5][(w >> 16) & 0xff] ^
        g_table[4][(w >> 24) & 0xff] ^ g_table[3][(w >> 32) & 0xff] ^ g_table[2][(w >> 40) & 0xff] ^
        g_table[1][(w >> 48) & 0xff] ^ g_table[0][(w >> 56) & 0xff];
    p += 8;
    n -= 8;
  }
  while (n--) c = g_table[0][(c ^ *p++) & 0xff] ^ (c >> 8);
  return c ^ 0xffffffffu;
}

// Scans a TFRecord buffer: fills offsets[i], lengths[i] of up to max_records payloads and returns
// the number found, or -(1 + index) of the first record whose length/data CRC does not verify.
extern "C" long long drn_tfrecord_index(const uint8_t* buf, size_t n, uint64_t* offsets,
                                        uint64_t* lengths, long long max_records, int verify) {
  size_t pos = 0;
  long long count = 0;
  while (pos + 12 <= n && count < max_records) {
    uint64_t len;
    memcpy(&len, buf + pos, 8);
    uint32_t len_crc;
    memcpy(&len_crc, buf + pos + 8, 4);
    if (verify) {
      uint32_t c = drn_crc32c(buf + pos, 8, 0);
      c = ((c >> 15) | (c << 17)) + 0xa282ead8u;
      if (c != len_crc) return -(1 + count);
    }
    if (pos + 12 + len + 4 > n) break;
    if (verify) {
      uint32_t data_crc;
      memcpy(&data_crc, buf + pos + 12 + len, 4);
      uint32_t c = drn_crc32c(buf + pos + 12, len, 0);
      c = ((c >> 15) | (c << 17)) + 0xa282ead8u;
      if (c != data_crc) return -(1 + count);
    }
    offsets[count] = pos + 12;
    lengths[count] = len;
    ++count;
    pos += 12 + len + 4;
  }
  return count;
}
