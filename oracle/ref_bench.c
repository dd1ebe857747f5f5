/* ref_bench.c -- threaded CPU baseline harness (TEST/BENCH INFRASTRUCTURE, ours).
 * Links the UNMODIFIED reference objects (oracle/_ref) and times opus_decode() on the host cores:
 * T threads, each owning n/T independent OpusDecoder objects, all decoding the same packet
 * sequence a GPU batch gets (stream i starts at packet (37*i) mod count of a .bit fixture).
 * Usage: ref_bench <file.bit> <Fs> <channels> <streams> <frames> <threads> [ms_streams ms_coupled mapping,csv]
 * With the three optional arguments the decoders are OpusMSDecoder objects of that layout
 * (opus_multistream_decode; zero-length records of the fixture are lost packets).
 * Decoder objects are created before the clock starts and destroyed after it stops: the timed region holds
 * decode calls only.
 * Prints one line: stream_seconds wall_seconds threads checksum
 */
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "opus.h"
#include "opus_multistream.h"

typedef struct { unsigned char *data; int len; } Packet;
static Packet *g_pk; static int g_count;
static int g_fs, g_ch, g_frames;
static int g_ms_streams, g_ms_coupled;
static unsigned char g_mapping[256];
typedef struct { int first, last; uint32_t sum; double audio; void **dec; } Job;
static pthread_barrier_t g_start;

static void *worker(void *arg) {
  Job *job = (Job *)arg;
  int err, i, t, k;
  opus_int16 *pcm = (opus_int16 *)malloc(sizeof(opus_int16) * 5760 * g_ch);
  uint32_t sum = 0;
  job->dec = (void **)calloc(job->last - job->first + 1, sizeof(void *));
  for (i = job->first; i < job->last; i++) {
    job->dec[i - job->first] = g_ms_streams
        ? (void *)opus_multistream_decoder_create(g_fs, g_ch, g_ms_streams, g_ms_coupled, g_mapping, &err)
        : (void *)opus_decoder_create(g_fs, g_ch, &err);
    if (!job->dec[i - job->first]) { fprintf(stderr, "create failed\n"); exit(1); }
  }
  pthread_barrier_wait(&g_start);          /* main takes t0 after this barrier */
  pthread_barrier_wait(&g_start);
  for (i = job->first; g_ms_streams && i < job->last; i++) {
    OpusMSDecoder *dec = (OpusMSDecoder *)job->dec[i - job->first];
    for (t = 0; t < g_frames; t++) {
      Packet *p = &g_pk[(37 * i + t) % g_count];
      int ret = opus_multistream_decode(dec, p->len ? p->data : NULL, p->len, pcm, p->len ? 5760 : g_fs / 50, 0);
      if (ret < 0) { fprintf(stderr, "decode failed %d\n", ret); exit(1); }
      for (k = 0; k < ret * g_ch; k += 97) sum = sum * 31u + (uint16_t)pcm[k];
      job->audio += (double)ret / g_fs;
    }
  }
  for (i = job->first; !g_ms_streams && i < job->last; i++) {
    OpusDecoder *dec = (OpusDecoder *)job->dec[i - job->first];
    for (t = 0; t < g_frames; t++) {
      Packet *p = &g_pk[(37 * i + t) % g_count];
      int ret = opus_decode(dec, p->data, p->len, pcm, 5760, 0);
      if (ret < 0) { fprintf(stderr, "decode failed %d\n", ret); exit(1); }
      for (k = 0; k < ret * g_ch; k += 97) sum = sum * 31u + (uint16_t)pcm[k];
      job->audio += (double)ret / g_fs;
    }
  }
  job->sum = sum;
  pthread_barrier_wait(&g_start);          /* main takes t1 after this barrier */
  for (i = job->first; i < job->last; i++) {
    if (g_ms_streams) opus_multistream_decoder_destroy((OpusMSDecoder *)job->dec[i - job->first]);
    else opus_decoder_destroy((OpusDecoder *)job->dec[i - job->first]);
  }
  free(job->dec);
  free(pcm);
  return NULL;
}

int main(int argc, char **argv) {
  if (argc < 7) { fprintf(stderr, "usage: %s file.bit Fs ch streams frames threads\n", argv[0]); return 2; }
  FILE *f = fopen(argv[1], "rb");
  if (!f) { perror(argv[1]); return 2; }
  g_fs = atoi(argv[2]); g_ch = atoi(argv[3]);
  int streams = atoi(argv[4]); g_frames = atoi(argv[5]);
  int T = atoi(argv[6]);
  if (argc >= 10) {
    g_ms_streams = atoi(argv[7]); g_ms_coupled = atoi(argv[8]);
    int c = 0;
    for (char *tok = strtok(argv[9], ","); tok && c < 255; tok = strtok(NULL, ",")) g_mapping[c++] = (unsigned char)atoi(tok);
  }
  int cap = 1024;
  g_pk = (Packet *)malloc(sizeof(Packet) * cap);
  for (;;) {
    unsigned char h[8];
    if (fread(h, 1, 8, f) != 8) break;
    int len = (h[0] << 24) | (h[1] << 16) | (h[2] << 8) | h[3];
    if (g_count == cap) { cap *= 2; g_pk = (Packet *)realloc(g_pk, sizeof(Packet) * cap); }
    g_pk[g_count].data = (unsigned char *)malloc(len + 1);
    g_pk[g_count].len = len;
    if (fread(g_pk[g_count].data, 1, len, f) != (size_t)len) break;
    g_count++;
  }
  fclose(f);
  if (T < 1) T = 1;
  if (T > streams) T = streams;
  pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * T);
  Job *jobs = (Job *)calloc(T, sizeof(Job));
  struct timespec t0, t1;
  pthread_barrier_init(&g_start, NULL, T + 1);
  for (int i = 0; i < T; i++) {
    jobs[i].first = (int)((long long)streams * i / T);
    jobs[i].last = (int)((long long)streams * (i + 1) / T);
    pthread_create(&th[i], NULL, worker, &jobs[i]);
  }
  pthread_barrier_wait(&g_start);          /* every decoder exists */
  clock_gettime(CLOCK_MONOTONIC, &t0);
  pthread_barrier_wait(&g_start);          /* go */
  pthread_barrier_wait(&g_start);          /* every frame decoded */
  clock_gettime(CLOCK_MONOTONIC, &t1);
  double audio = 0; uint32_t sum = 0;
  for (int i = 0; i < T; i++) { pthread_join(th[i], NULL); audio += jobs[i].audio; sum ^= jobs[i].sum; }
  double wall = (t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec);
  printf("%.6f %.6f %d %08x\n", audio, wall, T, sum);
  return 0;
}
