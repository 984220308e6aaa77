"""torch.autograd.Function wrappers around the libnfcore kernels.

The flow-VAE objective (optimization.py:66-92) is assembled by *user* closures (`encode`,
`decode`, `logp` in src/learning_mnist.py:57-99) that slice, reshape and transpose the
network outputs.  Those view operations are recorded by torch's tape (plumbing); every
arithmetic-heavy node is one of the Functions below, whose forward and backward are
libnfcore kernels.
"""
import torch

from . import ops


class MLP(torch.autograd.Function):
    """Feedforward.forward (nn_models.py:41-84): (weights [P], x [N, D_in]) -> out [N, D_out]."""

    @staticmethod
    def forward(ctx, weights, x, dims, hidden_act, out_act):
        weights = weights.contiguous()
        x = x.contiguous()
        out, ws = ops.mlp_forward(weights, x, dims, hidden_act, out_act)
        ctx.save_for_backward(weights, x, ws, out)
        ctx.cfg = (dims, hidden_act, out_act)
        return out

    @staticmethod
    def backward(ctx, g_out):
        weights, x, ws, out = ctx.saved_tensors
        dims, hidden_act, out_act = ctx.cfg
        need_gx = ctx.needs_input_grad[1]
        g_w, g_x = ops.mlp_backward(weights, x, dims, hidden_act, out_act, ws, out, g_out.contiguous(), need_gx)
        return g_w, g_x, None, None, None


class AmortisedFlow(torch.autograd.Function):
    """optimization.py:74-85: (z0, mu0, ls0, W, U, B) -> (zK [N, D], logdet [N])."""

    @staticmethod
    def forward(ctx, z0, mu0, ls0, W, U, Bb, log_eps):
        z0, mu0, ls0, W, U, Bb = (t.contiguous() for t in (z0, mu0, ls0, W, U, Bb))
        zK, logdet, tsave = ops.amortised_forward(z0, mu0, ls0, W, U, Bb, log_eps, save=True)
        ctx.save_for_backward(z0, ls0, W, U, zK, tsave)
        ctx.log_eps = log_eps
        return zK, logdet

    @staticmethod
    def backward(ctx, g_zK, g_logdet):
        z0, ls0, W, U, zK, tsave = ctx.saved_tensors
        N = z0.shape[0]
        if g_zK is None:
            g_zK = torch.zeros_like(zK)
        # dF/dlogdet[n] stays on the device (per-sample vector; optimization.py:85 averages, so it is -1/N everywhere)
        g_ld = 0.0 if g_logdet is None else g_logdet
        g_mu0, g_ls0, g_W, g_U, g_B = ops.amortised_backward(z0, ls0, W, U, zK, tsave, g_zK.contiguous(), g_ld, ctx.log_eps)
        return None, g_mu0, g_ls0, g_W, g_U, g_B, None


class BernoulliLL(torch.autograd.Function):
    """distributions.log_bern_mult (distributions.py:86-89): scalar over the whole batch."""

    @staticmethod
    def forward(ctx, x, p, log_eps):
        x = x.contiguous()
        p = p.contiguous()
        out, _ = ops.bernoulli_ll(x, p, log_eps, None)
        ctx.save_for_backward(x, p)
        ctx.log_eps = log_eps
        return out.reshape(())

    @staticmethod
    def backward(ctx, g):
        x, p = ctx.saved_tensors
        _, g_p = ops.bernoulli_ll(x, p, ctx.log_eps, g)          # the upstream scalar is read on the device
        return None, g_p, None
