"""BaseRNN with the reference's interface (reference: hmm_layer/BaseRNN.py:149-277): a time loop over a cell.

Kept for code that drives an HmmCell step by step (every step is one launch of the cell-step kernel).  The layer
functions in MsaHMMLayer.py do not use it: they run the whole scan as fused device passes.  Unlike the reference,
the tuple-state branch is chosen by type, not by len(hx) == 2 (defect D14)."""
import torch
import torch.nn as nn


class BaseRNN(nn.Module):
    def __init__(self, cell, batch_first=False, return_sequences=True, return_state=False, reverse=False):
        super().__init__()
        self.cell = cell
        self.batch_first = batch_first
        self.return_sequences = return_sequences
        self.return_state = return_state
        self.reverse = reverse

    def get_initial_state(self, inputs, batch_size):
        return self.cell.get_initial_state(inputs=inputs, batch_size=batch_size)

    def forward(self, inputs, hidden=None):
        """inputs (batch, seq, features) if batch_first else (seq, batch, features); hidden = cell state
        (for an HmmCell: [scaled, loglik]).  Returns per the return_sequences / return_state flags (reference :174-248)."""
        if self.batch_first:
            inputs = inputs.transpose(0, 1)
        if self.reverse:
            inputs = torch.flip(inputs, [0])
        seq_len, batch_size = inputs.shape[0], inputs.shape[1]
        hx = hidden if hidden is not None else self.get_initial_state(inputs, batch_size)
        tuple_state = isinstance(hx, (list, tuple))
        outputs = []
        for t in range(seq_len):
            if tuple_state:
                output, hx = self.cell(inputs[t], hx)
            else:
                hx = self.cell(inputs[t], hx)
                output = hx
            outputs.append(output)
        if self.return_sequences:
            out = torch.stack(outputs, dim=0)
            if self.batch_first:
                out = out.transpose(0, 1)
        else:
            out = outputs[-1]
            if self.batch_first:
                out = out.unsqueeze(1)
        if self.return_state:
            return out, hx
        return out
