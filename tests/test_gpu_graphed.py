"""GPU: a whole training pass (forward integration + adjoint backward) captured in a CUDA graph replays to the
same numbers as the eager pass, for new inputs too."""
import pytest
import torch

from oracle import graph as og
from test_gpu_partition import OPT

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('function,block', [('transformer', 'attention'), ('laplacian', 'attention'),
                                            ('laplacian', 'constant')])
def test_graphed_pass_equals_eager(function, block):
    import grand_b200
    n, d = 600, 64
    opt = dict(OPT, function=function, block=block, hidden_dim=d, attention_dim=32, step_size=0.5)
    torch.manual_seed(1)
    blk = grand_b200.set_block(opt)(grand_b200.set_function(opt), [], opt, torch.device('cuda'),
                                    t=torch.tensor([0., 2.])).cuda()
    blk.train()
    data = grand_b200.GraphData()
    data.new_graph(torch.tensor(og.shaped_graph(n, 5000, 2), device='cuda'), n)
    params = [p for p in blk.parameters() if p.requires_grad]

    def whole_pass(x_in, g_in):
        x = x_in.detach().requires_grad_(True)
        blk.set_x0(x)
        z = blk(x, data)
        loss = (z * g_in).sum()
        return (loss, z) + torch.autograd.grad(loss, [x] + params, allow_unused=True)

    gen = torch.Generator(device='cuda').manual_seed(0)
    x1, g1, x2, g2 = (torch.randn(n, d, device='cuda', generator=gen) for _ in range(4))
    graphed = grand_b200.GraphedPass(whole_pass, [x1, g1])
    for x, g in ((x1, g1), (x2, g2), (x1, g2)):
        want = [None if t is None else t.clone() for t in whole_pass(x, g)]
        got = graphed(x, g)
        torch.cuda.synchronize()
        for a, b in zip(got, want):
            assert (a is None) == (b is None)
            if a is not None:
                assert torch.equal(a, b)            # same kernels, same order: bit-identical
