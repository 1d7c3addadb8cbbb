import sys, os, time
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import numpy as np, torch, columba_b200 as cb, golden_util as gu
base=gu.index_of({}, None)
ix=cb.Index(base,4,10,4,device=0)
m=cb.Mapper(ix)
names,seqs=gu.read_fastq_gz(os.path.join(gu.GOLD,'g1_edit.fq.gz'))
seqs=[s for s in seqs if len(s)==150][:500]
reads,off=cb.Mapper.pack(seqs)
dr=torch.from_numpy(reads).cuda(); do=torch.from_numpy(off.view(np.int32)).cuda()
torch.cuda.synchronize()
t=time.time()
r=m.map_batch_device(dr.data_ptr(), do.data_ptr(), len(seqs), 150, 4, 0, fetch=False)
print("device call ok", r.counters['TOTAL_UNIQUE_MATCHES'], time.time()-t, flush=True)
