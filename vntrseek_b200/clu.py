"""Readers for the files on the PROCLU boundary (no computation): the counters block psearch
prints to stderr (src/psearch1.91/psearch.c:847-861) and the cluster file in a canonical form
(member order inside a `.clu` line follows heap addresses in the reference,
so files are compared as sets of clusters with relative directions)."""
import re

COUNTERS = ("CandidatesConsidered", "LCSalignments", "ProfileAlignments", "ProfileAlignmentsSuccess",
            "pairs confirmed after flank LCS alignments", "total clustered")


def parse_counters(log):
    out = {}
    for name in COUNTERS:
        m = re.search(r"\t" + re.escape(name) + r": (\d+)", log)
        out[name] = int(m.group(1)) if m else None
    return out


def canonical_clu(path):
    """clusters -> sorted list of ((key, dir), ...) with the smallest key normalised to dir 0"""
    clusters = []
    for line in open(path):
        toks = line.split()
        assert toks and toks[0] == "@", line
        mem = []
        for t in toks[1:]:
            mem.append((int(t[:-1]), {"'": 0, '"': 1, "+": 0, "-": 1}[t[-1]]))
        mem.sort()
        flip = mem[0][1]
        clusters.append(tuple((k, d ^ flip) for k, d in mem))
    clusters.sort()
    return clusters


def stats_as_counters(stats, trcount):
    """the six counters from a vsk_proclu_stats dict + vsk_clusters_write's trcount"""
    return {"CandidatesConsidered": stats["candidates_considered"], "LCSalignments": stats["lcs_alignments"],
            "ProfileAlignments": stats["profile_alignments"],
            "ProfileAlignmentsSuccess": stats["profile_alignments_success"],
            "pairs confirmed after flank LCS alignments": stats["confirmed"], "total clustered": trcount}
