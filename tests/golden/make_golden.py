"""Writes the golden fixtures under tests/golden/.

The reference is a Rust crate and there is no Rust toolchain in this image or on the GPU box, so the
reference itself cannot emit golden vectors.  Two kinds of fixture are committed instead:

  kat.json    the known-answer table of SURVEY.md §8c, verbatim: values derived by an independent numpy
              restatement that agrees with brute-force enumeration over internal states to 1e-15.  They pin
              the oracle (tests/test_oracle_kat.py, tests/test_golden.py) and, through it, the CUDA path.
  cases.json  seeded self-contained cases (Newick with f32 branch lengths, FASTA, model) with the outputs of
              the pinned oracle in reference-faithful mode: total lnL, per-pattern lnL, ASR state strings.
              They are regression vectors: the `-m gpu` tests compare the CUDA path with them WITHOUT the
              oracle in the loop, and the CPU tests check that the oracle still reproduces them.

Run from the repo root:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from oracle import RefAlignment, parse_newick, workloads  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

KAT = {
    "source": "SURVEY.md §8c known-answer table (independent numpy restatement + brute-force enumeration); "
              "branch lengths rounded to f32 as the reference's Newick parser does",
    "tiny_tree": "((A:0.601,B:0.301)C:0.100,D:0.200)E:0.001;",
    "aln4": ">A\nACGT\n>B\nAGGT\n>C\nACGT\n>D\nTTTT\n",
    "aln16": ">A\nACGTACGTACGTACGT\n>B\nAGGTAGGTAGGTAGGT\n>C\nACGTACGTACGTACGT\n>D\nTTTTTTTTTTTTTTTT\n",
    "aln8": ">A\nAAAAAAAA\n>B\nAAAAAAAC\n>C\nAAAAAAAA\n>D\nCCCCCCCC\n",
    "lnl": {
        "jc69_aln4": -16.577012400513222,
        "jc69_aln4_per_site": [-4.44798554852317, -5.200730091190169, -4.44798554852317, -2.480311212276715],
        "jc69_aln16": -66.3080496020529,
        "hky85_aln16": -68.11049472687087,
        "gtr_g4_aln16": -69.66427000171137,
        "gtr_i_g4_aln16": -71.7982617885503,
        "jc69_aln8": -35.0004062187667,
        "doc_example": -30.098779761644973,
        "n_minus_gap": -1.3862943611198906,
    },
    "doc_example": {"tree": "((A:0.1,B:0.2):0.15,(C:0.3,D:0.1):0.05);",
                    "aln": ">A\nACGTACGT\n>B\nACGTATGT\n>C\nACGAACGT\n>D\nTCGTACGA\n",
                    "model": "hky85(pi=1/4, kappa=2).with_gamma(0.5, 4)"},
    "discrete_gamma": {
        "0.5": [0.03338775338359941, 0.25191591759343773, 0.8202684819736465, 2.8944278470493163],
        "1.0": [0.1369537826446569, 0.4767518562354515, 0.999999999999997, 2.3862943611198943],
        "2.0": [0.2932747161235139, 0.6550136761581504, 1.0699896623909892, 1.9817219453273465],
    },
    "eigenvalues": {
        "hky85_pi.3.2.2.3_k2.5": [0.0, -0.9090909090909091, -1.5909090909090908, -1.5909090909090908],
        "gtr_pi.1.2.3.4_r123456": [0.0, -0.7042302112436655, -1.3512778335426976, -1.726543237264919],
    },
    "transition_row_A": {
        "hky85_pi.3.2.2.3_k2.5_t0.1": [0.9150977418327833, 0.01737985674354748, 0.0414526163083481, 0.026069785115321452],
        "gtr_pi.1.2.3.4_r123456_t0.25": [0.8533404980667257, 0.017548421334631817, 0.045295766825067696, 0.08381531377357469],
        "jc69_t0.5": [0.635062839274444, 0.121645720241852, 0.121645720241852, 0.121645720241852],
    },
    "asr": {
        "jc69_aln4_states": {"E": "ATGT", "C": "AGGT", "A": "ACGT", "B": "AGGT", "D": "TTTT"},
        "jc69_aln4_root_posterior_site0": [0.5011792911334836, 0.03100425268467755, 0.03100425268467753, 0.43681220349716127],
        "gtr_g4_aln16_states": {"E": "T" * 16, "C": "AGGT" * 4, "A": "ACGT" * 4, "B": "AGGT" * 4, "D": "T" * 16},
        "zero_branch_states_all": "T",
    },
}

CASES = [  # (name, n_taxa, n_sites, model, seed, ambiguity)
    ("jc69_12x60", 12, 60, "jc69", 101, 0.0),
    ("hky_g4_9x48", 9, 48, "hky_g4", 102, 0.0),
    ("gtr_g4_16x64_ambig", 16, 64, "gtr_g4", 103, 0.08),
    ("gtr_i_g4_24x80_ambig", 24, 80, "gtr_i_g4", 104, 0.05),
    ("gtr_i_7x40", 7, 40, "gtr_i", 105, 0.0),
]


def make_cases():
    out = []
    for name, n_taxa, n_sites, model, seed, amb in CASES:
        case = workloads.make_case(n_taxa, n_sites, model, seed=seed, ambiguity=amb)
        tree = case.tree
        newick = tree.to_newick()
        # the fixture is the TEXT form: ids are the ones the Newick parser assigns (newick.rs:168-284),
        # f32 lengths survive repr()
        tree2 = parse_newick(newick)
        assert tree2.to_newick() == newick
        rows = case.leaf_order
        fasta = "".join(f">{nm}\n{bytes(case.patterns[:, r]).decode()}\n" for r, nm in enumerate(rows))
        aln = RefAlignment.from_fasta_bytes(fasta.encode())
        pats, s2p, mult, lo = aln.compress_columns()
        ft = oracle.flatten(tree2, lo)
        tot, per = oracle.log_likelihood_flat(ft, case.model, pats, mult, mode=0)
        atot, _aper, states, post = oracle.marginal_asr_flat(ft, case.model, pats, mult, want_posteriors=True)
        assert abs(atot - tot) <= 1e-9 * abs(tot)
        srt = np.sort(post, axis=2)
        min_gap = float((srt[..., 3] - srt[..., 2]).min())
        seqs = {str(int(ft["ids"][k])): bytes(b"ACGT"[s] for s in states[k][s2p]).decode() for k in range(ft["n_nodes"])}
        # joint (Viterbi) ASR, compute_joint_asr likelihood.rs:421-560
        jtot, jper, jstates = oracle.joint_asr_flat(ft, case.model, pats, mult)
        jseqs = {str(int(ft["ids"][k])): bytes(b"ACGT"[s] for s in jstates[k][s2p]).decode() for k in range(ft["n_nodes"])}
        # NNI neighbours (caller loop clone + NNI::nni tree.rs:1213-1242 + log_likelihood), arena ids of the parsed tree
        nni = []
        for nid in tree2.get_node_ids():
            for lc in (False, True):
                t = tree2.clone()
                if t.nni(int(nid), lc):
                    continue
                val, _ = oracle.log_likelihood_flat(oracle.flatten(t, lo), case.model, pats, mult, mode=0)
                nni.append([int(nid), int(lc), float(val)])
        out.append(dict(name=name, model=model, newick=newick, fasta=fasta, log_likelihood=float(tot),
                        per_site=[float(x) for x in per[s2p]], asr_sequences=seqs, asr_min_top2_gap=min_gap,
                        joint_log_likelihood=float(jtot), joint_sequences=jseqs, nni_scores=nni,
                        n_patterns=int(len(pats)), site_to_pattern=[int(x) for x in s2p],
                        multiplicity=[int(x) for x in mult]))
    return out


if __name__ == "__main__":
    oracle.build()
    with open(os.path.join(HERE, "kat.json"), "w") as f:
        json.dump(KAT, f, indent=1)
    with open(os.path.join(HERE, "cases.json"), "w") as f:
        json.dump(dict(generator="tests/golden/make_golden.py (oracle, reference-faithful mode)", cases=make_cases()), f, indent=1)
    print("wrote kat.json, cases.json")
