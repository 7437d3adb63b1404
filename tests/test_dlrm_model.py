"""The plain-PyTorch DLRM module (CPU): single-index tables as in the reference notebook
(examples/pytorch_dlrm.ipynb cell 9) and the multi-hot extension (bags with padding)."""
import torch

from raydp_b200.models import DLRM


def test_single_index_forward_shape_and_spec():
    model = DLRM(13, [50, 60, 70], 32, bottom=(64, 32), top=(32, 1))
    out = model(torch.randn(9, 13), torch.randint(0, 50, (9, 3)))
    assert out.shape == (9, 1)
    spec = model.fused_spec()
    assert spec.bag_sizes is None and spec.index_columns == 3 and spec.pooling == "sum"


def test_bags_pool_and_ignore_padding():
    torch.manual_seed(0)
    model = DLRM(4, [10, 20], 8, bottom=(8,), top=(4, 1), sparse_grad=False, bag_sizes=[3, 1], pooling="mean")
    assert model.fused_spec().index_columns == 4
    dense = torch.randn(5, 4)
    idx = torch.tensor([[1, 2, 3, 0], [1, -1, -1, 5], [4, 4, -1, 7], [0, -1, 9, 1], [2, 2, 2, 2]])
    out = model(dense, idx)
    # replacing a padded slot by any row must not change the result; replacing a live slot must
    idx2 = idx.clone(); idx2[1, 1] = -1; idx2[1, 2] = -1
    assert torch.equal(out, model(dense, idx2))
    same_mean = idx.clone(); same_mean[1, 1] = 1          # bag {1, 1} has the same mean as {1}
    assert torch.allclose(out, model(dense, same_mean), atol=1e-6)
    other = idx.clone(); other[1, 1] = 6
    assert not torch.allclose(out[1], model(dense, other)[1])
    # gradients reach only the live rows
    out.sum().backward()
    g = model.embeddings[0].weight.grad
    assert g[5:9].abs().sum() == 0 and g[1].abs().sum() > 0
