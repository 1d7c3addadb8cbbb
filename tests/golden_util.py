"""Helpers shared by the tests: golden inputs, expected tuples, index fixture."""
import gzip
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
WORK = os.path.join(HERE, "_work")
ROOT = os.path.dirname(HERE)


def read_fastq_gz(path):
    names, seqs = [], []
    with gzip.open(path, "rt") as f:
        while True:
            h = f.readline()
            if not h:
                break
            s = f.readline().strip()
            f.readline()
            f.readline()
            names.append(h[1:].split()[0])
            seqs.append(s.upper())
    return names, seqs


def cases():
    return json.load(open(os.path.join(GOLD, "cases.json")))


def expected(case):
    rows = []
    with gzip.open(os.path.join(GOLD, "expected", case + ".tsv.gz"), "rt") as f:
        for line in f:
            a = line.rstrip("\n").split("\t")
            rows.append((a[0], int(a[1]), a[2], int(a[3]), int(a[4]), a[5]))
    ctr = json.load(open(os.path.join(GOLD, "expected", case + ".json")))
    return rows, ctr


def scheme_arg(params):
    kind, arg = params.get("scheme", ("named", "columba"))
    if kind != "named":
        arg = os.path.join(GOLD, "schemes", arg)
    return kind, arg


def golden_fasta(name="g1"):
    os.makedirs(WORK, exist_ok=True)
    fa = os.path.join(WORK, name + ".fa")
    if not os.path.exists(fa):
        with gzip.open(os.path.join(GOLD, name + ".fa.gz"), "rb") as i, open(fa + ".tmp", "wb") as o:
            o.write(i.read())
        os.replace(fa + ".tmp", fa)
    return fa


def reads_file(p):
    """FASTQ of a case: g1_<reads>.fq.gz, or <reads>.fq.gz for the sets with their own reference."""
    return os.path.join(GOLD, ("%s.fq.gz" % p["reads"]) if "fasta" in p else ("g1_%s.fq.gz" % p["reads"]))


def index_of(p, lib_path):
    """Index basename of a case's reference (built on first use by the product's host-side index writer)."""
    import columba_b200 as cb
    name = p.get("fasta", "g1")
    fa = golden_fasta(name)
    base = os.path.join(WORK, name)
    if not os.path.exists(base + ".rev.brt") or not os.path.exists(base + ".meta"):
        cb.build_index(fa, base, 4, lib_path=lib_path)
    return base


# counters of the reference's -v report -> names in our counter block
CTR_MAP = {"nodes": "NODE_COUNTER", "matches": "TOTAL_UNIQUE_MATCHES", "reported": "TOTAL_REPORTED_POSITIONS",
           "mapped": "MAPPED_READS", "reads": "NUMBER_OF_READS", "intext": "IN_TEXT_STARTED",
           "aborted": "ABORTED_IN_TEXT_VERIF", "immediate": "IMMEDIATE_SWITCH", "searches": "SEARCH_STARTED"}
