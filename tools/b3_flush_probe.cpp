// tools/b3_flush_probe.cpp -- how often does the three-phase band decoder flush its tape (cb_celt_bands3.h)?
// Host build of the kernel source with the B3_PROBE counters; decodes an opus_demo .bit file and prints the
// histogram of flushes per frame, how many were forced by a full tape, and the tape / bit sizes a frame needs.
// The tape size (CB_B3_BUF, CB_B3_BITWORDS) was chosen with it: every stream of config 2 flushes exactly twice,
// which is what lets the stage rendezvous of kernel A line up.
//   g++ -std=c++17 -O2 [-DBIG] -o /tmp/b3_probe tools/b3_flush_probe.cpp && /tmp/b3_probe tests/golden/c2_96k_cbr_st.bit 2
// (-DBIG: an unbounded tape, to read off what a frame would need)
#include <stdio.h>
int g_b3_flush = 0, g_b3_full = 0, g_b3_maxbuf = 0, g_b3_maxbits = 0, g_b3_needbuf = 0, g_b3_needbits = 0;
#ifdef BIG
#define CB_B3_BUF 4096
#define CB_B3_BITWORDS 256
#endif
#define B3_PROBE 1
#include "../tests/emu/emu_celt.cpp"
#include <vector>
int main(int argc, char** argv) {
  FILE* f = fopen(argv[1], "rb");
  int ch = atoi(argv[2]);
  void* d = emu_opus_create(48000, ch);
  std::vector<short> pcm(5760 * ch);
  int hist[32] = {0}, n = 0, fullsum = 0;
  unsigned char hdr[8];
  while (fread(hdr, 1, 8, f) == 8) {
    unsigned len = (hdr[0] << 24) | (hdr[1] << 16) | (hdr[2] << 8) | hdr[3];
    std::vector<unsigned char> p(len + 1);
    if (len && fread(p.data(), 1, len, f) != len) break;
    g_b3_flush = 0; g_b3_full = 0;
    emu_opus_decode(d, len ? p.data() : nullptr, len, pcm.data(), 5760, 0);
    hist[g_b3_flush < 31 ? g_b3_flush : 31]++; fullsum += g_b3_full; n++;
  }
  for (int i = 0; i < 32; i++) if (hist[i]) printf("flushes %d: %d frames\n", i, hist[i]);
  printf("frames %d, flushes caused by a full tape: %d, max buf %d, max bits %d, need buf %d bits %d\n", n, fullsum, g_b3_maxbuf, g_b3_maxbits, g_b3_needbuf, g_b3_needbits);
}
