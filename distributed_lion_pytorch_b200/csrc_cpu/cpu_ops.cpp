// cpu_ops.cpp — host-side codec and vote for the gloo / CPU path (BASELINE config 1).
//
// The wire format and the vote are integer work; doing them with ATen ops costs several full
// passes over byte-expanded temporaries (the reference even expands to int64:
// distributed_lion.py:75-77,84-88 and torch.mode at :38).  These loops touch each input byte once.
// The floating-point part of the step stays in ATen on the CPU so that it remains bit-identical
// to the reference's own ops (ATen's vectorised and scalar loops round bf16 differently; see
// tests/test_lion_api.py).
#include <stdint.h>
#include <string.h>

extern "C" {

// bools[n] (0/1 bytes) -> packed[(n+7)/8], LSB first, pad bits zero
void dlion_cpu_pack_bits(const uint8_t* bools, int64_t n, uint8_t* packed) {
  const int64_t full = n >> 3;
#pragma omp parallel for schedule(static)
  for (int64_t j = 0; j < full; ++j) {
    uint64_t w;
    memcpy(&w, bools + 8 * j, 8);                       // eight 0/1 bytes
    // gather bit 0 of every byte into one byte: multiply trick (bytes are 0 or 1)
    packed[j] = (uint8_t)((w * 0x0102040810204080ull) >> 56);
  }
  if (n & 7) {
    uint8_t b = 0;
    for (int64_t i = full * 8; i < n; ++i) b |= (uint8_t)((bools[i] & 1u) << (i & 7));
    packed[full] = b;
  }
}

// gathered[world][row_bytes] -> cnt[n] = number of live ranks whose bit is 1
void dlion_cpu_count_votes(const uint8_t* gathered, int world, int64_t row_bytes, uint32_t live_mask, int64_t n,
                           int16_t* cnt) {
  const int64_t nbytes = (n + 7) >> 3;
#pragma omp parallel for schedule(static)
  for (int64_t j = 0; j < nbytes; ++j) {
    int16_t c[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int r = 0; r < world; ++r) {
      if (!((live_mask >> r) & 1u)) continue;
      const uint8_t b = gathered[(int64_t)r * row_bytes + j];
      for (int k = 0; k < 8; ++k) c[k] += (b >> k) & 1;
    }
    const int64_t base = j << 3;
    const int lim = (int)((n - base) < 8 ? (n - base) : 8);
    for (int k = 0; k < lim; ++k) cnt[base + k] = c[k];
  }
}

int dlion_cpu_version(void) { return 100; }
}
