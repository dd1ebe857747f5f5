// Developer tool: replays .bit fixtures through the host build of the decoder sources (and runs the host
// build of the encoder on synthetic PCM) as a plain executable, so that it can run under AddressSanitizer
// (a Python process with the instrumented library preloaded does not get through the SILK code).
//   g++ -std=c++17 -O1 -g -fsanitize=address,undefined -o asan_replay tools/asan_replay.cpp
//   ./asan_replay dec <file.bit> <Fs> <channels> <frame_size> [corrupt]  |  ./asan_replay enc <Fs> <channels> <frames>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "../tests/emu/emu_celt.cpp"
#undef CB_HOST_EMU
namespace enc {
#define CB_HOST_EMU 1
}
extern "C" {
void* emu_enc_create(int Fs, int channels, int application);
int emu_enc_ctl(void* state, int request, int value, int* out);
int emu_enc_encode(void* state, const short* pcm, int frame_size, unsigned char* data, int max_bytes);
void emu_enc_destroy(void* s);
}

int main(int argc, char** argv) {
  if (argc >= 6 && !strcmp(argv[1], "dec")) {
    FILE* f = fopen(argv[2], "rb");
    if (!f) return 2;
    const int Fs = atoi(argv[3]), ch = atoi(argv[4]), fsz = atoi(argv[5]);
    const bool corrupt = argc > 6;
    void* d = emu_opus_create(Fs, ch);
    std::vector<short> pcm((size_t)5760 * ch);
    unsigned seed = 12345, calls = 0;
    for (;;) {
      unsigned char h[8];
      if (fread(h, 1, 8, f) != 8) break;
      const int len = (h[0] << 24) | (h[1] << 16) | (h[2] << 8) | h[3];
      // exact-size heap buffer: reads past the packet are caught
      unsigned char* p = (unsigned char*)malloc(len > 0 ? len : 1);
      if (len && fread(p, 1, len, f) != (size_t)len) break;
      int use = len;
      if (corrupt && len > 2) {
        seed = seed * 1664525u + 1013904223u;
        const unsigned u = (seed >> 16) % 10;
        if (u < 3) for (int k = 0; k < 4; k++) { seed = seed * 1664525u + 1013904223u; p[1 + (seed >> 8) % (len - 1)] = (unsigned char)(seed >> 24); }
        else if (u < 5) use = 1 + (seed >> 8) % (len - 1);
        else if (u < 6) use = 0;
      }
      unsigned char* q = (unsigned char*)malloc(use > 0 ? use : 1);
      memcpy(q, p, use);
      emu_opus_decode(d, use ? q : nullptr, use, pcm.data(), use ? fsz : (fsz < Fs / 50 ? fsz : Fs / 50), 0);
      free(q);
      free(p);
      calls++;
    }
    emu_opus_destroy(d);
    printf("decoded %u calls\n", calls);
    return 0;
  }
  if (argc >= 5 && !strcmp(argv[1], "enc")) {
    const int Fs = atoi(argv[2]), ch = atoi(argv[3]), frames = atoi(argv[4]);
    void* e = emu_enc_create(Fs, ch, 2051);
    emu_enc_ctl(e, 4002, 96000, nullptr);
    const int fs = Fs / 50;
    std::vector<short> pcm((size_t)fs * ch);
    unsigned char out[1500];
    unsigned seed = 7;
    long bytes = 0;
    for (int t = 0; t < frames; t++) {
      for (size_t i = 0; i < pcm.size(); i++) { seed = seed * 1664525u + 1013904223u; pcm[i] = (short)((int)(seed >> 16) % 20000 - 10000) * ((t / 10) % 2); }
      unsigned char* o = (unsigned char*)malloc(1275);
      const int n = emu_enc_encode(e, pcm.data(), fs, o, 1275);
      if (n > 0) { memcpy(out, o, n); bytes += n; }
      free(o);
    }
    emu_enc_destroy(e);
    printf("encoded %d frames, %ld bytes\n", frames, bytes);
    return 0;
  }
  return 2;
}
