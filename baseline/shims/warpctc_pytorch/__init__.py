import torch


class CTCLoss(torch.nn.Module):
    """warpctc_pytorch.CTCLoss(out[T,N,H] logits, labels_cpu, output_sizes, target_sizes) -> summed loss."""

    def forward(self, acts, labels, act_lens, label_lens):
        lp = acts.float().log_softmax(2)
        return torch.nn.functional.ctc_loss(lp, labels, act_lens, label_lens, blank=0, reduction='sum',
                                            zero_infinity=True).reshape(1)
