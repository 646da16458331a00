"""Generates the committed golden fixtures in this directory.  Runs ONLY in the build container, where the
reference checkout is mounted at /root/reference (it does not exist on the GPU box, and no test reads it).

  python tests/golden/make_golden.py

1. kmer_tf_golden.npz  – the reference's only known-answer data: the saved TensorFlow outputs in
   /root/reference/tests/test_tf.ipynb (cells 1 and 3), parsed from the notebook JSON.
2. ref_genepred.npz    – outputs of the reference's own PyTorch objects (GenePredHMMEmitter,
   GenePredMultiHMMTransitioner, HmmCell, BaseRNN, TotalProbabilityCell, _get_total_forward_from_chunks)
   on small seeded inputs.  The un-vendored `learnMSA` import (MsaHMMLayer.py:5) is stubbed; forward and
   reverse cells own separate transitioner instances (avoids the shared-`reverse`-flag defect D1); inputs are
   cloned before every emitter call (the emitter mutates them, D5) and contain no 'N' (the reference's second
   k-mer call is wrong at N positions, D5 — N handling is pinned by the TF golden instead).
3. ref_dense67.npz     – the reference's generic HmmCell + BaseRNN driven by a minimal dense plugin with the
   profile-HMM sparsity pattern (L=32, q=67), the config-1 shape family.
"""
import json
import os
import re
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def parse_tf_tensor(txt, name):
    m = re.search(re.escape(name) + r":?\s*tf\.Tensor\(\s*(\[.*?\]), shape=\(([^)]*)\)", txt, re.S)
    body, shape = m.group(1), tuple(int(x) for x in m.group(2).split(",") if x.strip())
    vals = np.array([float(x) for x in re.findall(r"-?\d+\.?\d*(?:e-?\d+)?", body)], dtype=np.float32)
    return vals.reshape(shape)


def kmer_golden():
    nb = json.load(open(os.path.join(REF, "tests", "test_tf.ipynb")))
    out1 = "".join("".join(o.get("text", [])) for o in nb["cells"][1]["outputs"])
    out3 = "".join("".join(o.get("text", [])) for o in nb["cells"][3]["outputs"])
    g = dict(
        sequences=np.eye(5, dtype=np.float32)[None],          # ACGTN one-hot, as in the cell source
        k_mers_left=parse_tf_tensor(out3, "k_mers_left"),
        k_mers_right=parse_tf_tensor(out3, "k_mers_right"),
        encoded_kmer_left=parse_tf_tensor(out3, "encoded_kmer_left"),      # encode_kmer_string("ACGN")
        encoded_kmer_right=parse_tf_tensor(out3, "encoded_kmer_right"),
        fill_triangular_lower=parse_tf_tensor(out1, "Lower Triangular Matrix:\n"),
        fill_triangular_upper=parse_tf_tensor(out1, "Upper Triangular Matrix:\n"),
    )
    assert g["k_mers_left"].shape == (1, 5, 16, 4) and g["encoded_kmer_right"].shape == (64, 4)
    np.savez_compressed(os.path.join(HERE, "kmer_tf_golden.npz"), **g)
    print("kmer_tf_golden.npz", {k: v.shape for k, v in g.items()})


def import_reference():
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "hmm_layer"))
    for name in ["learnMSA", "learnMSA.msa_hmm", "learnMSA.msa_hmm.Utility"]:
        sys.modules[name] = types.ModuleType(name)
    sys.modules["learnMSA.msa_hmm.Utility"].deserialize = lambda x: x


CODONS = dict(start_codons=[("ATG", 1.)], stop_codons=[("TAG", .34), ("TAA", .33), ("TGA", .33)],
              intron_begin_pattern=[("NGT", .99), ("NGC", .005), ("NAT", .005)],
              intron_end_pattern=[("AGN", .99), ("ACN", .01)])


def genepred_golden():
    import torch
    from hmm_layer.MsaHmmCell import HmmCell
    from hmm_layer.gene_pred_hmm_emitter import GenePredHMMEmitter
    from hmm_layer.gene_pred_hmm_transitioner import GenePredMultiHMMTransitioner
    from hmm_layer.BaseRNN import BaseRNN
    from hmm_layer.TotalProbabilityCell import TotalProbabilityCell
    from hmm_layer.MsaHMMLayer import _get_total_forward_from_chunks

    B, T, q = 3, 48, 15
    g = torch.Generator().manual_seed(1234)
    cls = torch.softmax(2 * torch.randn(1, B, T, 15, generator=g), -1)
    nuc_idx = torch.randint(0, 4, (1, B, T), generator=g)
    # plant a few biologically meaningful motifs so constrained states get non-zero emissions
    motif = {"ATG": 5, "GTA": 12, "TAG": 30, "AGC": 20, "TAA": 40}
    for s, pos in motif.items():
        for i, ch in enumerate(s):
            nuc_idx[0, 0, pos + i] = "ACGT".index(ch)
    nuc = torch.nn.functional.one_hot(nuc_idx, 5).float()
    inputs = torch.cat([cls, nuc], -1)
    end_hints = torch.rand(1, B, 2, q, generator=g)
    emis_kernel = torch.randn(1, 13, 15, generator=g)

    def make_cell(reverse):
        np.random.seed(0)
        em = GenePredHMMEmitter(**CODONS)
        em.build((1, B, T, 15))
        with torch.no_grad():
            em.emission_kernel.copy_(emis_kernel)
        tr = GenePredMultiHMMTransitioner(initial_exon_len=200, initial_intron_len=4500, initial_ir_len=10000)
        with torch.no_grad():
            tr.transition_kernel += 1e-3           # keep all 23 edges (defect D6 drops exact zeros)
        cell = HmmCell([q], 15, em, tr)
        if reverse:
            cell.reverse_direction()
        cell.recurrent_init()
        return cell

    out = {}
    with torch.no_grad():
        cell, rcell = make_cell(False), make_cell(True)
        em, tr = cell.emitter[0], cell.transitioner
        out.update(inputs=inputs.numpy(), end_hints=end_hints.numpy(), emission_kernel=emis_kernel.numpy(),
                   transition_kernel=tr.transition_kernel.numpy().copy(), transition_indices=tr.indices[:, 1:],
                   A=tr.A.numpy().copy(), pi=cell.init_dist.numpy()[0].copy(), B13=em.B.numpy().copy(),
                   codon_probs=em.codon_probs.numpy().copy())
        E = cell.emission_probs(inputs.clone())
        out["E"] = E.numpy().copy()
        out["E_hints"] = cell.emission_probs(inputs.clone(), end_hints=end_hints).numpy().copy()
        out["E_training"] = cell.emission_probs(inputs.clone(), training=True).numpy().copy()

        # sequential forward, control flow of MsaHMMLayer.py:247-272 with the real BaseRNN signature
        Ef = E.reshape(B, T, q)
        rnn = BaseRNN(cell, batch_first=True, return_sequences=True, return_state=True)
        st0 = cell.get_initial_state(batch_size=B, parallel_factor=1)
        f1, st1 = cell(Ef[:, 0], st0, init=True)
        fr, stT = rnn(Ef[:, 1:], st1)
        fwd = torch.cat([f1.unsqueeze(1), fr], 1)
        out["fwd_out"] = fwd.numpy().copy()                      # (B,T,16) = log alpha-hat ‖ running loglik
        out["loglik"] = stT[1].numpy().copy().reshape(1, B)

        # sequential backward, control flow of MsaHMMLayer.py:343-371
        rrnn = BaseRNN(rcell, batch_first=True, return_sequences=True, return_state=True, reverse=True)
        rs0 = rcell.get_initial_state(batch_size=B, parallel_factor=1)
        b1, rs1 = rcell(Ef[:, -1], rs0, init=True)
        br, _ = rrnn(Ef[:, :-1], rs1)
        bwd = torch.flip(torch.cat([b1.unsqueeze(1), br], 1), [1])
        out["bwd_out"] = bwd.numpy().copy()                      # (B,T,16) = log R ‖ loglik accumulated to the right

        # chunked conditional forward (parallel_factor = 4), MsaHMMLayer.py:254-262 + MsaHmmCell.py:121-142
        pf = 4
        Ec = E.reshape(B * pf, T // pf, q)
        cs0 = cell.get_initial_state(batch_size=B * pf, parallel_factor=pf)
        c1, cs1 = cell(Ec[:, 0], cs0, init=True)
        cr, _ = rnn(Ec[:, 1:], cs1)
        cfwd = torch.cat([c1.unsqueeze(1), cr], 1)
        out["chunk_fwd_out"] = cfwd.numpy().copy()               # (B*pf, T/pf, 240)
        tp = BaseRNN(TotalProbabilityCell(cell), batch_first=True, return_sequences=True, return_state=True)

        class _TP(torch.nn.Module):                              # adapter: BaseRNN needs an explicit initial state
            def __init__(s, r): super().__init__(); s.r = r
            def forward(s, x): return s.r(x, s.r.cell.get_initial_state(batch_size=x.shape[0], dtype=torch.float32))
        total, ll_c = _get_total_forward_from_chunks(cfwd, cell, _TP(tp), B, T, parallel_factor=pf)
        out["chunk_total_fwd"] = total.numpy().copy()            # (1,B,T,15) log alpha via the chunked route
        out["chunk_loglik"] = ll_c.numpy().copy()
    np.savez_compressed(os.path.join(HERE, "ref_genepred.npz"), **out)
    print("ref_genepred.npz", {k: np.asarray(v).shape for k, v in out.items()})


def multicopy_golden():
    """4. ref_multicopy.npz – the reference's emitters outside the 15-state single-copy case:
    GenePredHMMEmitter(num_copies = 2, 3) (gene_pred_hmm_emitter.py:231-263 with the per-copy repeat :254-257) and
    SimpleGenePredHMMEmitter (:93-121), with end hints and the training flag; plus the 7-state transition matrix.
    (The reference's A for k > 1 has all-zero rows as shipped -- exact-zero kernel entries are dropped, D6/D11 -- so
    the multi-copy transitioner is pinned by the oracle's edge list, not by reference output.)"""
    import torch
    from hmm_layer.gene_pred_hmm_emitter import GenePredHMMEmitter, SimpleGenePredHMMEmitter
    from hmm_layer.gene_pred_hmm_transitioner import SimpleGenePredHMMTransitioner

    B, T = 2, 40
    g = torch.Generator().manual_seed(77)
    cls = torch.softmax(2 * torch.randn(1, B, T, 15, generator=g), -1)
    nuc_idx = torch.randint(0, 4, (1, B, T), generator=g)
    for s_, pos in {"ATG": 3, "GTA": 9, "TAG": 20, "AGC": 14, "TAA": 30}.items():
        for i, ch in enumerate(s_):
            nuc_idx[0, 0, pos + i] = "ACGT".index(ch)
    x = torch.cat([cls, torch.nn.functional.one_hot(nuc_idx, 5).float()], -1)
    out = dict(inputs=x.numpy())
    with torch.no_grad():
        for nc in (2, 3):
            em = GenePredHMMEmitter(num_copies=nc, **CODONS)
            em.build((1, B, T, 15))
            ker = torch.randn(em.emission_kernel.shape, generator=g)
            em.emission_kernel.copy_(ker)
            em.recurrent_init()
            hints = torch.rand(1, B, 2, 1 + 14 * nc, generator=g)
            out[f"kernel{nc}"] = ker.numpy()
            out[f"hints{nc}"] = hints.numpy()
            out[f"E{nc}"] = em(x.clone()).numpy().copy()
            out[f"E{nc}_hints_training"] = em(x.clone(), end_hints=hints, training=True).numpy().copy()
        x7 = torch.softmax(torch.randn(1, B, T, 7, generator=g), -1)
        sem = SimpleGenePredHMMEmitter()
        sem.build((1, B, T, 7))
        ker7 = torch.randn(sem.emission_kernel.shape, generator=g)
        sem.emission_kernel.copy_(ker7)
        sem.recurrent_init()
        hints7 = torch.rand(1, B, 2, 7, generator=g)
        st = SimpleGenePredHMMTransitioner()
        st.transition_kernel += 1e-3
        st.recurrent_init()
        out.update(inputs7=x7.numpy(), kernel7=ker7.numpy(), hints7=hints7.numpy(), E7=sem(x7.clone()).numpy().copy(),
                   E7_hints=sem(x7.clone(), end_hints=hints7).numpy().copy(), A7=st.A.numpy().copy(),
                   transition_kernel7=st.transition_kernel.numpy().copy(), pi7=st.make_initial_distribution().numpy()[0].copy())
    np.savez_compressed(os.path.join(HERE, "ref_multicopy.npz"), **out)
    print("ref_multicopy.npz", {k: np.asarray(v).shape for k, v in out.items()})


def dense_golden():
    import torch
    from hmm_layer.MsaHmmCell import HmmCell
    from hmm_layer.BaseRNN import BaseRNN
    from hmm_layer.Transitioner import _make_sparse_transition_indices_implicit

    L, B, T, s = 32, 2, 30, 24
    q = 2 * L + 3
    g = torch.Generator().manual_seed(2000)
    idx = np.concatenate([np.asarray(v).reshape(-1, 2) for v in _make_sparse_transition_indices_implicit(L).values()])
    A = torch.zeros(q, q)
    A[idx[:, 0], idx[:, 1]] = torch.rand(len(idx), generator=g) + 0.05
    A = (A / A.sum(-1, keepdim=True)).unsqueeze(0)
    pi = torch.softmax(torch.randn(1, 1, q, generator=g), -1)
    Bm = torch.softmax(torch.randn(1, q, s, generator=g), -1)
    x = torch.nn.functional.one_hot(torch.randint(0, s, (1, B, T), generator=g), s).float()

    class DenseTransitioner(torch.nn.Module):
        reverse = False
        def recurrent_init(self): pass
        def make_log_A(self): return torch.log(A + 1e-30)
        def make_initial_distribution(self): return pi
        def forward(self, v): return torch.matmul(v, A.transpose(1, 2) if self.reverse else A)

    class DenseEmitter(torch.nn.Module):
        def recurrent_init(self): pass
        def forward(self, inputs, end_hints=None, training=False):
            return torch.einsum("kbts,kqs->kbtq", inputs, Bm)

    with torch.no_grad():
        cell = HmmCell([q], s, DenseEmitter(), DenseTransitioner())
        E = cell.emission_probs(x).reshape(B, T, q)
        rnn = BaseRNN(cell, batch_first=True, return_sequences=True, return_state=True)
        st0 = cell.get_initial_state(batch_size=B, parallel_factor=1)
        f1, st1 = cell(E[:, 0], st0, init=True)
        fr, stT = rnn(E[:, 1:], st1)
        fwd = torch.cat([f1.unsqueeze(1), fr], 1)
    out = dict(edges=idx, A=A.numpy(), pi=pi.numpy()[0], B=Bm.numpy(), inputs=x.numpy(), E=E.numpy().reshape(1, B, T, q),
               fwd_out=fwd.numpy(), loglik=stT[1].numpy().reshape(1, B))
    np.savez_compressed(os.path.join(HERE, "ref_dense67.npz"), **out)
    print("ref_dense67.npz", {k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "multicopy":     # only the fixture added in round 2
        import_reference()
        multicopy_golden()
        sys.exit(0)
    kmer_golden()
    import_reference()
    genepred_golden()
    dense_golden()
