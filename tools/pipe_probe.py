"""Developer probe: repeat the pipelined decode loop of bench.py several times in one process and print the
rate of each repetition (is the copy/kernel overlap stable?)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import concentus_b200 as cb
from bench import read_bit
lib = cb.load()
n, SPF, CH = 4096, 960, 2
pk = [p for p, _ in read_bit(os.path.join(ROOT, "tests/golden/c2_96k_cbr_st.bit"))]
plen = len(pk[0]); npk = len(pk)
table = np.stack([np.frombuffer(p, dtype=np.uint8) for p in pk])
rot = (37 * np.arange(n)) % npk
arenas = [torch.from_numpy(np.ascontiguousarray(table[(rot + t) % npk]).reshape(-1)).pin_memory() for t in range(64)]
offs = torch.from_numpy((np.arange(n, dtype=np.int32) * plen).astype(np.int32)).pin_memory()
lens = torch.from_numpy(np.full(n, plen, dtype=np.int32)).pin_memory()
bd = cb.OpusBatchDecoder(48000, CH, n)
h_pcm = [torch.zeros(n * SPF * CH, dtype=torch.int16).pin_memory() for _ in range(3)]
h_ret = [torch.zeros(n, dtype=torch.int32).pin_memory() for _ in range(3)]
t = 0
for rep in range(8):
    K = 100
    torch.cuda.synchronize()
    e0 = time.perf_counter()
    for k in range(K):
        a = arenas[t % 64]
        rc = lib.opus_decode_group_submit(bd._g, a.data_ptr(), a.numel(), offs.data_ptr(), lens.data_ptr(),
                                          h_pcm[k % 3].data_ptr(), SPF * CH, SPF, None, h_ret[k % 3].data_ptr())
        assert rc == 0
        t += 1
    assert lib.opus_decode_group_wait(bd._g) == 0
    e1 = time.perf_counter()
    print("rep %d: %.3f ms/step  %.0f stream-s/s" % (rep, 1e3 * (e1 - e0) / K, n * 0.02 * K / (e1 - e0)), flush=True)
bd.close()
