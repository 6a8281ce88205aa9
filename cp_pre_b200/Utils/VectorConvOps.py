"""Drop-in for ``Utils/VectorConvOps.py``: vector calculus on [BS, Nt, Nx, Ny] fields.

REFERENCE DEFECT, not reproduced by default: every constructor of the reference passes
``requires_grad`` positionally into ``ConvOperator``'s ``conv`` slot and therefore raises
``ValueError("Unknown Convolution Method")``.  A drop-in that always raises
is useless, so the children are built the way the reference's own test intends
(Tests/NS_vector_convops.py:154-176).  Set ``STRICT_REFERENCE = True`` to get the shipped
behaviour.  Divergence and Curl are single fused passes (one kernel, no temporaries).
"""
from cp_pre_b200 import _lib as _L
from cp_pre_b200 import ops as _ops
from cp_pre_b200.Utils.ConvOps_2d import *  # noqa: F401,F403
from cp_pre_b200.Utils.ConvOps_2d import ConvOperator

STRICT_REFERENCE = False


def dot(a, b):
    return a[0] * b[0] + a[1] * b[1]


def cross(a, b):
    return a[0] * b[1] + a[1] * b[0]      # plus sign, as shipped


def vectorize(a, b):
    return torch.stack((a, b))


def _child(domain, order, scale, taylor_order, requires_grad):
    if STRICT_REFERENCE:
        return ConvOperator(domain, order, scale, taylor_order, requires_grad)   # raises, like the reference
    return ConvOperator(domain, order, scale, taylor_order, requires_grad=requires_grad)


def _pair(op_x, op_y, a, b, sign):
    """D_x(a) + sign * D_y(b) in one fused pass when both children take the direct path."""
    if op_x.conv != op_x.convolution or op_y.conv != op_y.convolution:
        return op_x(a) + sign * op_y(b)
    staged = _ops.Staged(a)
    fa, _, _ = op_x._prep_field(a)
    fb, _, _ = op_y._prep_field(b)
    kx, ky = op_x._taps(op_x.kernel), op_y._taps(op_y.kernel)
    pads = [tuple(s // 2 for s in k.shape) for k in (kx, ky)]
    out = _ops.run_program(_L.PROG_NS_CONTINUITY, [_ops.to_cuda(fa), _ops.to_cuda(fb)], [kx, ky], pads,
                           [float(sign)])
    return staged.back(out)


class Divergence(ConvOperator):
    def __init__(self, domain=('x', 'y'), order=1, scale=1.0, taylor_order=2, require_grad=False):
        super(Divergence, self).__init__()
        self.grad_x = _child(domain[0], order, scale, taylor_order, require_grad)
        self.grad_y = _child(domain[1], order, scale, taylor_order, require_grad)

    def __call__(self, input_x, input_y):
        return _pair(self.grad_x, self.grad_y, input_x, input_y, 1.0)


class Gradient(ConvOperator):
    def __init__(self, domain=('x', 'y'), order=1, scale=1.0, taylor_order=2, requires_grad=False):
        super(Gradient, self).__init__()
        self.grad_x = _child(domain[0], order, scale, taylor_order, requires_grad)
        self.grad_y = _child(domain[1], order, scale, taylor_order, requires_grad)

    def __call__(self, input_x, input_y=None):
        if input_y is None:
            input_y = input_x
        return torch.stack((self.grad_x(input_x), self.grad_y(input_y)))


class Curl(ConvOperator):
    def __init__(self, domain=('x', 'y'), order=1, scale=1.0, taylor_order=2, requires_grad=False):
        super(Curl, self).__init__()
        self.grad_x = _child(domain[0], order, scale, taylor_order, requires_grad)
        self.grad_y = _child(domain[1], order, scale, taylor_order, requires_grad)

    def __call__(self, input_x, input_y):
        return _pair(self.grad_x, self.grad_y, input_y, input_x, -1.0)


class Laplace(ConvOperator):
    def __init__(self, domain=('x', 'y'), order=2, scale=1.0, taylor_order=2, requires_grad=False):
        super(Laplace, self).__init__()
        self.laplace = _child(domain, order, scale, taylor_order, requires_grad)

    def __call__(self, input_x, input_y=None):
        if input_y is None:
            input_y = input_x
        return torch.stack((self.laplace(input_x), self.laplace(input_y)))
