import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
from test_multistream import _sub_packets, LAYOUTS, pack_multistream
from mousiki_b200.multistream import MultistreamBatchDecoder
layout=LAYOUTS[0][1]
n_dec,ticks=6,7
subs=[_sub_packets(layout,ticks,seed=11+d) for d in range(n_dec)]
lost=np.random.default_rng(3).random((n_dec,ticks))<0.2
lost[:,0]=False
packets=[[None if lost[d,t] else pack_multistream(subs[d][t]) for t in range(ticks)] for d in range(n_dec)]
print(lost.astype(int))
for d in range(n_dec): print(d, [hex(subs[d][t][0][0]) for t in range(ticks)])
with MultistreamBatchDecoder(n_dec,48000,2,1,1,[0,1]) as ms:
    for t in range(ticks):
        print("tick",t,flush=True)
        pcm,s=ms.decode_float([packets[d][t] for d in range(n_dec)],960)
        print(s,flush=True)
