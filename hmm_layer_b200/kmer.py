"""k-mer helpers with the reference's names (reference: hmm_layer/kmer.py).

On the hot path the per-position soft 3-mers are never built: the CUDA emission stage reads the five
nucleotide flags around a position and looks the codon tables up directly (csrc/gene15.cu, codon_side).
These torch functions exist for API parity and for building the small (2,9,64) codon tables; unlike the
reference's torch port (kmer.py:23-25) they do not modify their input."""
import torch


def make_k_mers(sequences, k, pivot_left=True):
    """(b, L, 5) one-hot over A,C,G,T,N -> (b, L, 4**(k-1), 4) soft k-mers (reference kmer.py:3-47)."""
    p = sequences[..., :4] + 0.25 * (sequences[..., 4:5] == 1).to(sequences.dtype)
    b, L, _ = p.shape
    pad = torch.full((b, k - 1, 4), 0.25, dtype=p.dtype, device=p.device)
    if pivot_left:
        padded = torch.cat([p, pad], 1)
        base, offsets = padded[:, :L], range(1, k)
    else:
        padded = torch.cat([pad, p], 1)
        base, offsets = padded[:, k - 1:L + k - 1], range(k - 2, -1, -1)
    out = base.unsqueeze(2)
    for i in offsets:
        shift = padded[:, i:L + i]
        out = (out.unsqueeze(3) * shift[:, :, None, :, None]).reshape(b, L, -1, 4)
    return out


def encode_kmer_string(kmer, pivot_left=True, alphabet="ACGT"):
    """(4**(k-1), 4) encoding of a k-mer string over ACGTN (reference kmer.py:49-65)."""
    full = alphabet + "N"
    idx = torch.tensor([full.index(c) for c in kmer])
    one_hot = torch.nn.functional.one_hot(idx, len(full)).to(torch.float32)
    enc = make_k_mers(one_hot.unsqueeze(0), k=len(kmer), pivot_left=pivot_left)[0]
    return enc[0] if pivot_left else enc[-1]
