// tests/emu/emu_celt.cpp -- TEST-ONLY build of the kernel bodies as plain C++ (CB_HOST_EMU):
// every lane loop becomes an ordinary loop, every warp fold an identity.  It lets the CPU-side
// tests (-m "not gpu") check the codec logic of csrc/*.h against the oracle without a GPU.
// It is NOT part of the product: libopus_b200.so never links it and has no CPU path.
#define CB_HOST_EMU 1
#include <stdlib.h>
#include <string.h>
#include "../../concentus_b200/csrc/cb_celt_entropy.h"
#include "../../concentus_b200/csrc/cb_celt_synth.h"

extern "C" {

void* emu_celt_create(int channels, int downsample) {
  CeltDecState* s = (CeltDecState*)malloc(sizeof(CeltDecState));
  celt_state_reset(s, channels, downsample);
  return s;
}
void emu_celt_destroy(void* s) { free(s); }

// Decode one CELT frame (payload without TOC).  Returns samples per channel or <0.
int emu_celt_decode(void* state, const unsigned char* payload, int len, int LM, int C, int start, int end,
                    short* pcm, unsigned* final_range) {
  CeltDecState* st = (CeltDecState*)state;
  static EntWork ew;
  static SynWork sw;
  static FrameSide side;
  EntFrameCfg cfg = {LM, C, start, end, 0};
  int rc = celt_entropy_decode(ew, st, &side, payload, len, cfg, final_range);
  if (rc < 0) return rc;
  for (int c = 0; c < st->channels; c++) celt_synth_channel(sw, st, &side, c, pcm);
  return (120 << LM) / st->downsample;
}

}  // extern "C"
