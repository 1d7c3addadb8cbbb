"""N > 1 host logic on CPU: two gloo ranks shard the golden reads, map their
batches (through the host emulation of the kernels -- test harness, no GPU
here) and reduce counters and per-read hit counts; the totals must equal the
reference's golden output of the unsharded run."""
import os
import socket
import sys

import numpy as np
import pytest

import golden_util as gu

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, base, lib, out_dir):
    sys.path.insert(0, gu.ROOT)
    sys.path.insert(0, HERE)
    import torch.distributed as dist
    import columba_b200 as cb
    from columba_b200 import parallel as par
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    ix = cb.Index(base, 4, 10, 4, lib_path=lib)
    m = cb.Mapper(ix)
    names, seqs = gu.read_fastq_gz(os.path.join(gu.GOLD, "g1_edit.fq.gz"))
    sm = par.ShardedMapper(m, rank, world, batch=300)
    results, ctr, st, hits = sm.map_reads(seqs, 4)
    tup = []
    for b, r in results:
        tup += r.tuples(names[b:b + r.n_reads], ix.seq_names)
    gathered = [None] * world
    dist.all_gather_object(gathered, tup)
    if rank == 0:
        allt = sorted(t for g in gathered for t in g)
        np.save(os.path.join(out_dir, "hits.npy"), hits)
        import json
        json.dump({"ctr": ctr, "tuples": allt}, open(os.path.join(out_dir, "out.json"), "w"))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_batches_cover_input_once():
    from columba_b200 import parallel as par
    for n, world, batch in ((0, 2, 10), (1, 2, 10), (95, 2, 10), (100, 4, 10), (1000, 8, 64)):
        seen = np.zeros(n, int)
        for r in range(world):
            for b, e in par.shard_batches(n, r, world, batch):
                seen[b:e] += 1
        assert (seen == 1).all()


@pytest.mark.timeout(600)
def test_two_ranks_equal_unsharded_golden(golden_index, hostemu_lib, tmp_path):
    import json
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_rank_main, args=(2, port, golden_index, hostemu_lib, str(tmp_path)), nprocs=2, join=True)
    out = json.load(open(tmp_path / "out.json"))
    exp, ctr = gu.expected("edit_k4_columba")
    got = [tuple(t) for t in out["tuples"]]
    assert got == exp
    for k, v in gu.CTR_MAP.items():
        assert out["ctr"][v] == ctr[k], k
    hits = np.load(tmp_path / "hits.npy")
    names, _ = gu.read_fastq_gz(os.path.join(gu.GOLD, "g1_edit.fq.gz"))
    per = {}
    for t in exp:
        per[t[0]] = per.get(t[0], 0) + 1
    assert [int(h) for h in hits] == [per.get(nm, 0) for nm in names]
