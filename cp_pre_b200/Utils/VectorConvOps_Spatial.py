"""Drop-in for ``Utils/VectorConvOps_Spatial.py``: vector calculus on [BS, Nvar, Nx, Ny] fields.

Each operator pads its inputs with ``BoundaryManager((taylor+1, taylor+1))`` and applies the
valid stencil of its child ``ConvOps_Spatial.ConvOperator``; here padding and correlation are
one libcppre kernel reading through the boundary index map (no padded copies).  Kept from the
reference: ``grad_x`` and ``grad_y`` hold identical kernels (both differentiate along dim 2);
``Gradient`` / ``Vector_Gradient`` build their children on ``cuda`` with ``requires_grad``.
"""
from cp_pre_b200 import ops as _ops
from cp_pre_b200.Utils.ConvOps_Spatial import *  # noqa: F401,F403
from cp_pre_b200.Utils.ConvOps_Spatial import ConvOperator
from cp_pre_b200.Utils.boundary_conditions import *  # noqa: F401,F403
from cp_pre_b200.Utils.boundary_conditions import BoundaryManager


def dot(a, b):
    return a[:, 0:1] * b[:, 0:1] + a[:, 1:2] * b[:, 1:2]


def cross(a, b):
    # plus sign, as shipped [ref: Utils/VectorConvOps_Spatial.py:26-27]
    return a[:, 0:1] * b[:, 1:2] + a[:, 1:2] * b[:, 0:1]


def vectorize(a, b):
    return torch.cat((a, b), dim=1)


class _BCOperator(ConvOperator):
    """Shared: boundary manager + 'pad then valid-correlate with a child's live kernel'."""

    def _setup_bc(self, taylor_order, boundary_cond):
        self.bc = BoundaryManager(kernel_size=(taylor_order + 1, taylor_order + 1))
        self.bc.set_all_boundaries(bc_type=boundary_cond)

    def _apply(self, child, x):
        if child.conv != child.convolution:              # spectral children: pad, then the child
            return child(self.bc.pad_signal(x))
        staged = _ops.Staged(x)
        xc = _ops.to_cuda(x)
        if xc.dim() != 4 or xc.shape[1] != 1:
            raise RuntimeError(f"expected input [BS, 1, Nx, Ny], got {tuple(x.shape)}")
        taps = child._taps(child.kernel)
        return staged.back(_ops.bc_conv2d(xc, self.bc.descriptor(), taps))


class Gradient(_BCOperator):
    # 1 -> 2
    def __init__(self, domain=('x', 'y'), order=1, scale=1.0, taylor_order=2, boundary_cond='periodic',
                 conv='direct', device=torch.device("cpu"), requires_grad=False):
        super(Gradient, self).__init__()
        self.grad_x = ConvOperator(domain[0], order, scale, taylor_order, conv, device=torch.device("cuda"),
                                   requires_grad=True)
        self.grad_y = ConvOperator(domain[1], order, scale, taylor_order, conv, device=torch.device("cuda"),
                                   requires_grad=True)
        self._setup_bc(taylor_order, boundary_cond)

    def __call__(self, input_x, input_y=None):
        if input_y is None:
            input_y = input_x
        return torch.cat((self._apply(self.grad_x, input_x), self._apply(self.grad_y, input_y)), dim=1)


class Laplace(_BCOperator):
    # scalar: 1 -> 1, vector: 2 -> 2
    def __init__(self, domain=('x', 'y'), order=2, scale=1.0, taylor_order=2, boundary_cond='periodic',
                 scalar=True, conv='direct', device='cpu', requires_grad=False):
        super(Laplace, self).__init__()
        self.laplace = ConvOperator(domain, order, scale, taylor_order, conv, device, requires_grad)
        self.scalar = scalar
        self._setup_bc(taylor_order, boundary_cond)

    def __call__(self, input_x, input_y=None):
        if self.scalar == True:  # noqa: E712
            return self._apply(self.laplace, input_x)
        if input_y is None:
            input_y = input_x
        return torch.cat((self._apply(self.laplace, input_x), self._apply(self.laplace, input_y)), dim=1)


class Divergence(_BCOperator):
    # 2 -> 1
    def __init__(self, domain=('x', 'y'), order=1, scale=1.0, taylor_order=2, boundary_cond='periodic',
                 conv='direct', device='cpu', requires_grad=False):
        super(Divergence, self).__init__()
        self.grad_x = ConvOperator(domain[0], order, scale, taylor_order, conv, device, requires_grad)
        self.grad_y = ConvOperator(domain[1], order, scale, taylor_order, conv, device, requires_grad)
        self._setup_bc(taylor_order, boundary_cond)

    def __call__(self, input_x, input_y):
        return self._apply(self.grad_x, input_x) + self._apply(self.grad_y, input_y)


class Curl(_BCOperator):
    # 2 -> 1
    def __init__(self, domain=('x', 'y'), order=1, scale=1.0, taylor_order=2, boundary_cond='periodic',
                 conv='direct', device='cpu', requires_grad=False):
        super(Curl, self).__init__()
        self.grad_x = ConvOperator(domain[0], order, scale, taylor_order, conv, device, requires_grad)
        self.grad_y = ConvOperator(domain[1], order, scale, taylor_order, conv, device, requires_grad)
        self._setup_bc(taylor_order, boundary_cond)

    def __call__(self, input_x, input_y):
        return self._apply(self.grad_x, input_y) - self._apply(self.grad_y, input_x)


class Vector_Gradient(_BCOperator):
    # 2 -> 1:  gx(a)^2 + gy(b)^2 + 2 gy(a) gx(b)
    def __init__(self, domain=('x', 'y'), order=1, scale=1.0, taylor_order=2, boundary_cond='periodic',
                 conv='direct', device=torch.device("cpu"), requires_grad=False):
        super(Vector_Gradient, self).__init__()
        self.grad_x = ConvOperator(domain[0], order, scale, taylor_order, conv, device=torch.device("cuda"),
                                   requires_grad=True)
        self.grad_y = ConvOperator(domain[1], order, scale, taylor_order, conv, device=torch.device("cuda"),
                                   requires_grad=True)
        self._setup_bc(taylor_order, boundary_cond)

    def __call__(self, input_x, input_y):
        gxa, gyb = self._apply(self.grad_x, input_x), self._apply(self.grad_y, input_y)
        return gxa ** 2 + gyb ** 2 + 2 * self._apply(self.grad_y, input_x) * self._apply(self.grad_x, input_y)
