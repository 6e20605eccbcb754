"""ORACLE (test infrastructure): BASELINE.json config 1 -- the reference's bundled data/Over-Sampled_Tiny_Trans-c.csv -- as a
small committed fixture (tests/golden/config1_tiny.npz), generated in the build container:

    python -m oracle.gen_config1_fixture

Follows the reference's own preparation: rows read as src/datasets/ibm_transactions_for_aml.py:225-239 does
(``header=0``: the file has no header line, its first row is consumed as one -> 5 198 transactions), account ids made
unique per bank and compacted in first-appearance order as data/prepare_AML_transactions.py:24-35, columns typed as
ibm_transactions_for_aml.py:240-248 (3 categorical, 1 numerical, 1 timestamp -> K = 5), label = 'Is Laundering'.
Only the table is stored (ids, raw column values, category codes); no reference source is copied.
"""
import os

import numpy as np
import pandas as pd

from .ref_loader import REF_ROOT

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "config1_tiny.npz")
NAMES = ['Timestamp', 'From Bank', 'From ID', 'To Bank', 'To ID', 'Amount Received', 'Receiving Currency', 'Amount Paid',
         'Payment Currency', 'Payment Format', 'Is Laundering']


def main():
    df = pd.read_csv(os.path.join(REF_ROOT, "data", "Over-Sampled_Tiny_Trans-c.csv"), names=NAMES, header=0)
    src = df['From ID'].astype(str) + df['From Bank'].astype(str)
    dst = df['To ID'].astype(str) + df['To Bank'].astype(str)
    uniq = pd.concat([src, dst]).unique()
    idmap = {old: new for new, old in enumerate(uniq)}
    edge_index = np.stack([src.map(idmap).to_numpy(np.int64), dst.map(idmap).to_numpy(np.int64)])
    t = pd.to_datetime(df['Timestamp'], unit='s')
    ts7 = np.stack([t.dt.year, t.dt.month - 1, t.dt.day - 1, t.dt.weekday, t.dt.hour, t.dt.minute, t.dt.second], 1).astype(np.int64)
    cats, ncat = [], []
    for c in ('Receiving Currency', 'Payment Currency', 'Payment Format'):      # category order: frequency, as StatType.COUNT lists it
        order = df[c].value_counts().index.tolist()
        cats.append(df[c].map({v: i for i, v in enumerate(order)}).to_numpy(np.int64))
        ncat.append(len(order))
    np.savez_compressed(OUT, edge_index=edge_index, ts7=ts7[:, None, :], amount=df['Amount Paid'].to_numpy(np.float32)[:, None],
                        cat=np.stack(cats, 1), ncat=np.asarray(ncat), y=df['Is Laundering'].to_numpy(np.int64),
                        num_nodes=np.asarray(len(uniq)))
    print(f"config1_tiny: E = {edge_index.shape[1]}, N = {len(uniq)}, positives = {int(df['Is Laundering'].sum())}, ncat = {ncat}")


if __name__ == "__main__":
    main()
