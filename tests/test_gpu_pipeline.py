"""HostLabelPipeline (host-fed stage loop of generate_pseudo_label.py:60-96 + cal_prototype.py:63-82 +
segmentation_evaluate.py:38-44): five streams, two slots, label kernel ahead of the batch, co-resident kernels, tail
stream.  Whatever the overlap, the results must be those of the plain calls issued one after the other on one stream:
thresholded labels and fp16 confidences bit-equal, the confusion matrix exact, the prototype bank equal up to the order
of the fp32 segmented sums (2e-6) with identical counts.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from b200seg import dropin, ops, pipeline
    return dropin, ops, pipeline


def _batches(n_batches, n, c, k, h, w, H, W):
    g = torch.Generator().manual_seed(4711)
    out = []
    for _ in range(n_batches):
        logits = (3.0 * torch.randn(n, c, h, w, generator=g)).pin_memory()
        feat = torch.randn(n, k, h, w, generator=g).pin_memory()
        gt = torch.randint(0, c + 1, (n, H, W), generator=g).to(torch.uint8).pin_memory()   # c = not a class
        out.append((logits, feat, gt))
    return out


@pytest.mark.parametrize("overlap", [True, False])
def test_pipeline_matches_the_plain_calls(mods, overlap):
    dropin, ops, pipeline = mods
    n, c, k, h, w, H, W = 4, 13, 256, 41, 81, 320, 640
    thr, keep, ignore = 0.3, (5, 11), 12
    batches = _batches(5, n, c, k, h, w, H, W)
    pipe = pipeline.HostLabelPipeline("cuda", c, k, (H, W), thr=thr, keep=keep, ignore_index=ignore, overlap=overlap)
    # the same work, one call after the other on the current stream
    ref_feat = dropin.Class_Features("cuda", c, k)
    ref_hist = torch.zeros(c, c, dtype=torch.int64, device="cuda")
    got_labels, got_conf = [], []
    for i, (logits, feat, gt) in enumerate(batches):
        assert pipe.submit(logits, feat, gt) == i
        pipe.wait(i)                                        # slot i % 2 is rewritten by batch i + 2: read it now
        got_labels.append(pipe.out_label[i % 2].clone())
        got_conf.append(pipe.out_conf[i % 2].clone())
    pipe.join()
    torch.cuda.synchronize()
    for i, (logits, feat, gt) in enumerate(batches):
        dl, df, dg = logits.cuda(), feat.cuda(), gt.cuda()
        pl = ops.pseudo_label(dl, size=(H, W), thr=thr, keep=keep, ignore_index=ignore)
        assert torch.equal(got_labels[i], pl.label_thr.cpu()), "thresholded labels of batch %d" % i
        assert torch.equal(got_conf[i].view(torch.int16), pl.conf16.cpu().view(torch.int16)), "confidences of batch %d" % i
        ops.confusion_matrix(pl.label.reshape(-1), dg.reshape(-1), c, ref_hist)
        ops.class_means(df, dl, c, update=(ref_feat.objective_vectors, ref_feat.objective_vectors_num, ops.UPDATE_MEAN, True,
                                           ref_feat.proto_momentum))
    torch.cuda.synchronize()
    assert torch.equal(pipe.meter.hist, ref_hist)
    assert int(ref_hist.sum()) == sum(int((b[2] < c).sum()) for b in batches)
    assert torch.equal(pipe.features.objective_vectors_num, ref_feat.objective_vectors_num)
    a, b = pipe.features.objective_vectors.double().cpu().numpy(), ref_feat.objective_vectors.double().cpu().numpy()
    assert np.abs(b).max() > 0
    assert np.abs(a - b).max() <= 2e-6 * np.abs(b).max()
    assert pipe.h2d_bytes == sum(t.numel() * t.element_size() for t in batches[0])
    assert pipe.d2h_bytes == n * H * W * 3
