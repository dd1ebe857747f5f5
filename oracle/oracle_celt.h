/* oracle_celt.h -- interface of the plain-C oracle (TEST INFRASTRUCTURE; see oracle_celt.c). */
#ifndef ORACLE_CELT_H
#define ORACLE_CELT_H
#include <stdint.h>
struct OracleCelt;
struct OracleCelt *oracle_celt_create(int channels, int downsample);
void oracle_celt_destroy(struct OracleCelt *st);
void oracle_celt_reset(struct OracleCelt *st);
void oracle_celt_set_bands(struct OracleCelt *st, int start, int end);
int oracle_celt_pitch(const struct OracleCelt *st);
/* One CELT frame (payload without TOC); data==NULL or len<=1 conceals.  frame_size48 = samples at 48 kHz.
 * C = channels coded in the stream.  Returns samples per channel at the decoder's rate, or <0. */
int oracle_celt_decode(struct OracleCelt *st, const uint8_t *data, int len, int16_t *pcm, int frame_size48, int C,
                       int accum, uint32_t *final_range);

/* Opus layer (oracle_opus.c) */
struct OracleOpus;
struct OracleOpus *oracle_opus_create(int Fs, int channels);
void oracle_opus_destroy(struct OracleOpus *st);
int oracle_opus_decode(struct OracleOpus *st, const uint8_t *data, int len, int16_t *pcm, int frame_size, int decode_fec);
uint32_t oracle_opus_final_range(const struct OracleOpus *st);
void oracle_opus_set_gain(struct OracleOpus *st, int gain);
struct OracleMS;
struct OracleMS *oracle_ms_create(int Fs, int channels, int streams, int coupled, const uint8_t *mapping);
void oracle_ms_destroy(struct OracleMS *st);
int oracle_ms_decode(struct OracleMS *st, const uint8_t *data, int len, int16_t *pcm, int frame_size, int decode_fec);
uint32_t oracle_ms_final_range(const struct OracleMS *st);
#endif
