#!/usr/bin/env python
"""Extract the hard-coded search-scheme TABLES (pi/L/U strings, seeding
positions, weights, static begins) of the reference's built-in strategies
(src/searchstrategy.h:2829-3736) into the search_schemes/ directory format
(further_info/advanced_options/README.md:36-134) under
columba_b200/schemes/<name>/.  Data only; run once in the build container.
"""
import os
import re
import sys

src = open("/root/reference/src/searchstrategy.h").read()
out_root = sys.argv[1] if len(sys.argv) > 1 else "columba_b200/schemes"

CLASSES = {
    "KucherovKPlus1": "kuch1", "KucherovKPlus2": "kuch2",
    "OptimalKianfar": "kianfar", "O1StarSearchStrategy": "01star0",
    "PigeonHoleSearchStrategy": "pigeon", "MinUSearchStrategy": "minU",
    "ColumbaSearchStrategy": "columba_high", "DynamicColumbaStrategy": "columba_mid",
}

def class_body(name):
    m = re.search(r"^class %s\b.*?^};" % name, src, re.S | re.M)
    return m.group(0)

def searches(block):
    res = []
    for m in re.finditer(r"makeSearch\(\s*\{([^}]*)\}\s*,\s*\{([^}]*)\}\s*,\s*\{([^}]*)\}\s*,\s*(\d+)\s*\)", block, re.S):
        f = lambda s: "{" + ",".join(x.strip() for x in s.split(",")) + "}"
        res.append("%s %s %s" % (f(m.group(1)), f(m.group(2)), f(m.group(3))))
    return res

def vecvec(body, var):
    m = re.search(var + r"\s*=\s*\{(.*?)\};", body, re.S)
    if not m:
        return None
    return [[x.strip() for x in v.split(",") if x.strip()] for v in re.findall(r"\{([^{}]*)\}", m.group(1))]

for cls, name in CLASSES.items():
    body = class_body(cls)
    d = os.path.join(out_root, name)
    os.makedirs(d, exist_ok=True)
    open(os.path.join(d, "name.txt"), "w").write(name + "\n")
    if cls == "DynamicColumbaStrategy":
        for k in (2, 4, 6):
            m = re.search(r"getMidSearch%d\(\)\s*\{(.*?)\n    \}" % k, body, re.S)
            os.makedirs(os.path.join(d, str(k)), exist_ok=True)
            open(os.path.join(d, str(k), "searches.txt"), "w").write("\n".join(searches(m.group(1))) + "\n")
        continue
    seeds = vecvec(body, "seedingPositions")
    weights = vecvec(body, "weights")
    begins = vecvec(body, "staticPositions")
    for m in re.finditer(r"const std::vector<Search> ED(\d+) = \{(.*?)\};", body, re.S):
        k = int(m.group(1))
        kd = os.path.join(d, str(k))
        os.makedirs(kd, exist_ok=True)
        open(os.path.join(kd, "searches.txt"), "w").write("\n".join(searches(m.group(2))) + "\n")
        idx = k - 1
        if seeds is not None and weights is not None and idx < len(seeds):
            open(os.path.join(kd, "dynamic_partitioning.txt"), "w").write(
                " ".join(seeds[idx]) + "\n" + " ".join(weights[idx]) + "\n")
        if begins is not None and idx < len(begins):
            open(os.path.join(kd, "static_partitioning.txt"), "w").write(" ".join(begins[idx]) + "\n")
print("ok")
