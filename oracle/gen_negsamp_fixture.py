"""Writes tests/golden/negsamp_fixture.npz from the reference's own fixture graph
(/root/reference/src/primitives/negative_sampling/{edge_index,pos_edge_index}.json: one B = 200 HI-Small
batch, E_b = 10 702, 12 797 nodes) plus one output of the COMPILED reference sampler (oracle/_ref) on it."""
import json, os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "_ref"))
import negative_sampling as ref   # noqa: E402
D = "/root/reference/src/primitives/negative_sampling"
ei = json.load(open(f"{D}/edge_index.json")); pos = json.load(open(f"{D}/pos_edge_index.json"))
neg = ref.generate_negative_samples(ei, pos, 64)
np.savez_compressed(os.path.join(HERE, "..", "tests", "golden", "negsamp_fixture.npz"),
                    edge_index=np.asarray(ei, np.int64), pos_edge_index=np.asarray(pos, np.int64), ref_neg=np.asarray(neg, np.int64))
print("written", np.asarray(neg).shape)
