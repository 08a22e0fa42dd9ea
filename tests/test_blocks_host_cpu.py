"""CPU: the host-side arithmetic of MixedODEblock / HardAttODEblock (torch glue executed once per block forward)
against the reference's formulas (src/block_mixed.py:29-33, src/block_transformer_hard_attention.py:33-56),
and the block / function registry."""
import pytest
import torch


def test_registry_maps_the_reference_keys():
    import grand_b200
    from grand_b200 import model_configurations as mc
    assert mc.set_block({'block': 'constant'}) is grand_b200.ConstantODEblock
    assert mc.set_block({'block': 'attention'}) is grand_b200.AttODEblock
    assert mc.set_block({'block': 'mixed'}) is grand_b200.MixedODEblock
    assert mc.set_block({'block': 'hard_attention'}) is grand_b200.HardAttODEblock
    assert mc.set_function({'function': 'laplacian'}) is grand_b200.LaplacianODEFunc
    assert mc.set_function({'function': 'transformer'}) is grand_b200.ODEFuncTransformerAtt
    with pytest.raises(NotImplementedError):
        mc.set_block({'block': 'rewire_attention'})
    assert mc.set_function({'function': 'GAT'}) is grand_b200.ODEFuncAtt
    with pytest.raises(mc.BlockNotDefined):
        mc.set_block({'block': 'nope'})
    with pytest.raises(mc.FunctionNotDefined):
        mc.set_function({'function': 'nope'})


@pytest.mark.parametrize('n', [1, 2, 7, 1000, 4097])
@pytest.mark.parametrize('q', [0.0, 0.1, 0.5, 0.75, 1.0])
def test_linear_quantile_is_torch_quantile(n, q):
    from grand_b200.block_transformer_hard_attention import linear_quantile
    v = torch.rand(n, generator=torch.Generator().manual_seed(n))
    torch.testing.assert_close(linear_quantile(v, q), torch.quantile(v, q), rtol=1e-6, atol=1e-7)


def test_hard_attention_sampling_follows_the_reference():
    from grand_b200.block_transformer_hard_attention import hard_attention_mask, renormalise_attention
    g = torch.Generator().manual_seed(0)
    n, e, heads, pct = 30, 400, 4, 0.6
    ei = torch.randint(0, n, (2, e), generator=g)
    att = torch.rand(e, heads, generator=g)
    mean_att = att.mean(dim=1)
    mask = hard_attention_mask(mean_att, pct)
    ref_mask = mean_att > torch.quantile(mean_att, 1 - pct)                          # reference :54-55
    assert torch.equal(mask, ref_mask)
    assert abs(int(mask.sum()) - pct * e) <= 1
    kept = ei[:, mask]
    for idx in (0, 1):
        got = renormalise_attention(mean_att[mask], kept[idx], n)
        sums = torch.zeros(n).scatter_add_(0, kept[idx], mean_att[mask])             # torch_scatter.scatter(reduce='sum')
        torch.testing.assert_close(got, mean_att[mask] / (sums[kept[idx]] + 1e-16))
        total = torch.zeros(n).index_add_(0, kept[idx], got)
        present = torch.zeros(n, dtype=torch.bool)
        present[kept[idx]] = True
        torch.testing.assert_close(total[present], torch.ones(int(present.sum())))   # every kept group sums to 1


def test_mixed_attention_formula_and_gate_gradient():
    from grand_b200.block_mixed import mix_attention
    g = torch.Generator().manual_seed(1)
    att = torch.rand(1, 50, 4, generator=g)
    ew = torch.rand(1, 50, generator=g)
    gamma = torch.nn.Parameter(0.3 * torch.ones(1))
    mixed = mix_attention(att, ew, gamma)
    s = torch.sigmoid(gamma.detach())
    torch.testing.assert_close(mixed, att.mean(dim=2) * (1 - s) + ew * s)            # reference :29-33
    mixed.sum().backward()
    torch.testing.assert_close(gamma.grad, ((ew - att.mean(dim=2)).sum() * s * (1 - s)).reshape(1))
    assert float(mix_attention(att, ew, torch.tensor(-40.0)).sub(att.mean(dim=2)).abs().max()) < 1e-6
