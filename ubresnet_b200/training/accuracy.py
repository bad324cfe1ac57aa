"""`accuracy(output, target)` of the reference training scripts (training/train_ubresnet2018_wlarcv2.py:509-566; the
five-value variant with the track+shower "non-zero" accuracy: training/train_ubresnet2018_wlarcv1.py:528-586).

The reference computes argmax, a correctness mask and 2*C masked sums with one `.item()` host sync each.  Here one
kernel (`ubr_accuracy_counts`) produces all 2*C integer counters and ONE device->host copy reads them back.  Unlike the
reference, `target` is not resized in place (:528 mutates it); it only needs B*H*W elements.
"""
from .. import ops


def accuracy_counts(output, target):
    """device int64 tensor (2*C,): pixels per class, then correctly predicted pixels per class."""
    return ops.accuracy_counts(output.detach(), target)


def accuracy(output, target, imgdata=None, nonzero=None):
    """[acc_class0 .. acc_class{C-1}, total acc] in percent (0.0 for an absent class), as the wlarcv2 script; with
    `imgdata` given (wlarcv1 signature) or nonzero=True a sixth value -- accuracy over classes 1 and 2 -- is appended."""
    ncls = output.size(1)
    cnt = accuracy_counts(output, target).tolist()            # the only host sync
    num, corr = cnt[:ncls], cnt[ncls:]
    res = [100.0 * corr[c] / num[c] if num[c] > 0 else 0.0 for c in range(ncls)]
    res.append(100.0 * float(sum(corr)) / float(sum(num)))
    if nonzero or (nonzero is None and imgdata is not None):
        res.append(100.0 * float(corr[1] + corr[2]) / float(num[1] + num[2]))
    return res
