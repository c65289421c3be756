"""ag2_classify_images_static (no counterpart in the reference):
the LeNet pass with its activations in a fixed device array inside libag2.so, for callers
that may not allocate.  Checked through the C ABI against ag2_classify_images, the existing entry
point, on two calls in a row.  The kernels and the split-K choice are the same, so the logits must be
the same bits; the tolerance of the other LeNet tests (1e-3 absolute, 1e-5 relative) is the stated
bound.  Collected last (file name) so that a failure here hides no other test under pytest -x."""
import numpy as np
import pytest

from agile_grasp2_b200 import capi
from common import make_params
from oracle import binding as ob

pytestmark = pytest.mark.gpu
MAX_IMAGES = 64  # AG2_CLASSIFY_STATIC_MAX_IMAGES, include/ag2.h


def _images(seed, n):
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, size=(n, 60, 60, 3), dtype=np.uint8)
    img[rng.random(img.shape) < 0.6] = 0
    return img


def test_static_scratch_classifier_equals_the_allocating_one_twice():
    import torch
    w = ob.xavier_weights(23, bias_scale=0.2)
    g = capi.Context(make_params(capi.MODE_PREDICTION))
    g.set_lenet_weights(w)
    dev = torch.device("cuda", 0)
    # first call: a full batch; second call: fewer, different images (the array keeps the first call's
    # activations in the rows the second does not use), then the full batch again
    for seed, n in ((1, MAX_IMAGES), (2, 19), (1, MAX_IMAGES), (3, 1)):
        img = _images(seed, n)
        want = g.classify_images(img)
        d_img = torch.from_numpy(img).to(dev)
        d_log = torch.full((n, 2), float("nan"), dtype=torch.float32, device=dev)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        g.classify_images_static_ptr(d_img.data_ptr(), n, d_log.data_ptr())
        g.synchronize()
        assert torch.cuda.memory_allocated() == before
        got = d_log.cpu().numpy()
        tol = np.maximum(1e-3, 1e-5 * np.abs(want))
        assert np.all(np.abs(got - want) <= tol), (n, np.abs(got - want).max())
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), n  # same kernels, same split: same bits


def test_static_scratch_classifier_rejects_more_than_its_maximum():
    import torch
    g = capi.Context(make_params(capi.MODE_PREDICTION))
    g.set_lenet_weights(ob.xavier_weights(23, bias_scale=0.2))
    n = MAX_IMAGES + 1
    d_img = torch.zeros((n, 60, 60, 3), dtype=torch.uint8, device="cuda:0")
    d_log = torch.full((n, 2), 7.0, dtype=torch.float32, device="cuda:0")
    with pytest.raises(capi.Ag2Error) as e:
        g.classify_images_static_ptr(d_img.data_ptr(), n, d_log.data_ptr())
    assert e.value.code == capi.AG2_E_CAPACITY
    assert "AG2_CLASSIFY_STATIC_MAX_IMAGES" in str(e.value)
    g.synchronize()
    assert bool((d_log == 7.0).all())  # nothing was launched


def test_static_scratch_classifier_needs_weights():
    import torch
    g = capi.Context(make_params(capi.MODE_PREDICTION))
    d_img = torch.zeros((1, 60, 60, 3), dtype=torch.uint8, device="cuda:0")
    d_log = torch.zeros((1, 2), dtype=torch.float32, device="cuda:0")
    with pytest.raises(capi.Ag2Error) as e:
        g.classify_images_static_ptr(d_img.data_ptr(), 1, d_log.data_ptr())
    assert e.value.code == capi.AG2_E_NO_WEIGHTS
