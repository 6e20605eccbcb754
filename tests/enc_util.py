"""Synthetic transaction table of the IBM AML shape (src/datasets/ibm_transactions_for_aml.py: two amounts, two
currencies / one payment format, one timestamp -> K = 5 columns) for the column-encoder tests."""
import torch

from b200mp.nn import (EmbeddingEncoder, LinearEncoder, StatType, StypeWiseFeatureEncoder, TensorFrame, TimestampEncoder,
                       stype)


def make_frame(rows, seed=0, ncat=(15, 7), n_num=2, n_ts=1, missing=True, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    num = torch.randn(rows, n_num, generator=g) * 50 + 120
    cat = torch.stack([torch.randint(-1 if missing else 0, n, (rows,), generator=g) for n in ncat], 1)
    ts = torch.stack([torch.stack([torch.randint(2019, 2024, (rows,), generator=g), torch.randint(0, 12, (rows,), generator=g),
                                   torch.randint(0, 31, (rows,), generator=g), torch.randint(0, 7, (rows,), generator=g),
                                   torch.randint(0, 24, (rows,), generator=g), torch.randint(0, 60, (rows,), generator=g),
                                   torch.randint(0, 60, (rows,), generator=g)], 1) for _ in range(n_ts)], 1)
    names = {stype.numerical: [f"amount{i}" for i in range(n_num)], stype.categorical: [f"cat{i}" for i in range(len(ncat))],
             stype.timestamp: [f"time{i}" for i in range(n_ts)]}
    stats = {}
    for i, n in enumerate(names[stype.numerical]):
        stats[n] = {StatType.MEAN: float(num[:, i].mean()), StatType.STD: float(num[:, i].std()) if rows > 1 else 1.0}
    for i, n in enumerate(names[stype.categorical]):
        stats[n] = {StatType.COUNT: (list(range(ncat[i])), [1] * ncat[i])}
    for i, n in enumerate(names[stype.timestamp]):
        stats[n] = {StatType.YEAR_RANGE: [2019, 2023]}
    tf = TensorFrame({stype.numerical: num, stype.categorical: cat, stype.timestamp: ts}, names,
                     y=torch.randint(0, 2, (rows,), generator=g)).to(device)
    return tf, stats, names


def make_encoder(C, stats, names, seed=1):
    torch.manual_seed(seed)
    enc = StypeWiseFeatureEncoder(C, stats, names, {stype.categorical: EmbeddingEncoder(), stype.numerical: LinearEncoder(),
                                                   stype.timestamp: TimestampEncoder()})
    with torch.no_grad():                 # biases start at zero: give every parameter a value that shows up in the output
        for p in enc.parameters():
            if float(p.abs().max()) == 0:
                p.normal_(std=0.05)
    return enc


def oracle_blocks(enc, tf, dtype=torch.float64):
    """the encoder's parameters as oracle/stype_encoders.py blocks (leaf tensors of ``dtype`` requiring grad)"""
    e = enc.encoder_dict
    leaf = lambda t: t.detach().cpu().to(dtype).clone().requires_grad_(True)
    P = {"num_w": leaf(e["numerical"].weight), "num_b": leaf(e["numerical"].bias),
         "ts_w": leaf(e["timestamp"].weight), "ts_b": leaf(e["timestamp"].bias),
         "tabs": [leaf(m.weight) for m in e["categorical"].embs]}
    blocks = []
    for st in tf.stype_list:
        f = tf.feat_dict[st].cpu()
        if st is stype.numerical:
            blocks.append(("numerical", f.to(dtype), e["numerical"].mean.cpu().to(dtype), e["numerical"].std.cpu().to(dtype),
                           P["num_w"], P["num_b"]))
        elif st is stype.categorical:
            blocks.append(("categorical", f, P["tabs"]))
        else:
            blocks.append(("timestamp", f, e["timestamp"].min_year.cpu(), P["ts_w"], P["ts_b"]))
    return blocks, P
