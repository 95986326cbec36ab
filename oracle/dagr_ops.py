"""oracle/dagr_ops.py -- torch-CPU restatement of the third-party ops on the DAGR hot path.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU
baseline legs, never by eventad_b200/.

The reference calls these from libraries that are not vendored under /root/reference
(pins: requirements.txt:50-53): torch_spline_conv 1.2.2, torch_cluster 1.6.3,
torch_scatter 2.1.2, torch_geometric 2.6.1.  Their published algorithms are restated here
 next to the reference call site each one serves.  The reference has no golden
vectors for them -> **parity unpinned** for these functions (hand-derived known-answer tests in
tests/test_oracle_ops.py are the only anchor).
"""
import torch


# ---------------------------------------------------------------------------------------------
# torch_spline_conv.spline_basis  (call sites: src/dagr/model/layers/spline_conv.py:6,32 and PyG
# SplineConv.message).  Degree-1 basis:  k_mod=0 -> 1-v, k_mod=1 -> v, written as the library does:
#   1 - v - k_mod + 2*v*k_mod   (fp32, left to right)
# ---------------------------------------------------------------------------------------------
def spline_basis(pseudo: torch.Tensor, kernel_size: torch.Tensor, is_open_spline: torch.Tensor, degree: int = 1):
    assert degree == 1, "the path only uses degree 1 (spline_conv.py:10)"
    E, D = pseudo.shape
    S = (degree + 1) ** D
    basis = torch.empty((E, S), dtype=pseudo.dtype)
    index = torch.empty((E, S), dtype=torch.long)
    for s in range(S):
        k = s
        b = torch.ones(E, dtype=pseudo.dtype)
        wi = torch.zeros(E, dtype=torch.long)
        off = 1
        for d in range(D):
            k_mod = k % (degree + 1)
            k //= (degree + 1)
            ks = int(kernel_size[d])
            v = pseudo[:, d] * float(ks - degree * int(is_open_spline[d]))
            vi = v.to(torch.long)                              # (int64_t) v : truncation toward zero
            wi = wi + torch.fmod(vi + k_mod, ks) * off         # C++ % : remainder with the sign of the dividend
            off *= ks
            v = v - torch.floor(v)
            v = 1 - v - k_mod + 2 * v * k_mod
            b = b * v
        basis[:, s] = b
        index[:, s] = wi
    return basis, index


# torch_spline_conv.spline_weighting: out[e,o] = sum_s sum_i basis[e,s] * weight[index[e,s], i, o] * x[e,i]
def spline_weighting(x, weight, basis, index):
    out = torch.zeros((x.shape[0], weight.shape[2]), dtype=x.dtype)
    for s in range(basis.shape[1]):
        out += basis[:, s:s + 1] * torch.einsum("ni,nio->no", x, weight[index[:, s]])
    return out


# ---------------------------------------------------------------------------------------------
# torch_geometric.transforms.Cartesian(norm=True, max_value=m, cat=False)
# (call sites: src/dagr/model/layers/components.py:25-35, src/dagr/model/networks/net.py:70-93;
#  the async code restates the same formula: src/dagr/asynchronous/cartesian.py:6-16)
# ---------------------------------------------------------------------------------------------
def cartesian(pos: torch.Tensor, edge_index: torch.Tensor, max_value):
    row, col = edge_index[0], edge_index[1]
    cart = pos[row] - pos[col]
    if cart.numel() > 0:
        cart = cart * cuda_scalar_reciprocal(2 * max_value) + 0.5
    return cart


def cuda_scalar_reciprocal(scalar):
    """`tensor / python_scalar` as torch evaluates it ON CUDA, where the reference runs: the scalar is converted to fp32,
    its reciprocal is taken in fp32 and the tensor is MULTIPLIED by it (ATen/native/cuda/BinaryDivTrueKernel.cu,
    `iter.is_cpu_scalar(2)` branch); differs from the CPU's true division by up to 1 ulp."""
    return torch.tensor(1.0, dtype=torch.float32) / torch.tensor(float(scalar), dtype=torch.float32)


# ---------------------------------------------------------------------------------------------
# torch_cluster.grid_cluster(pos, size, start, end)  (call site: pooling.py:56)
# ---------------------------------------------------------------------------------------------
def grid_cluster(pos, size, start, end):
    pos = pos.view(pos.size(0), -1)
    pos = pos - start.unsqueeze(0)
    num_voxels = ((end - start) / size).to(torch.long) + 1
    num_voxels = num_voxels.cumprod(0)
    num_voxels = torch.cat([torch.ones(1, dtype=torch.long), num_voxels], 0)
    num_voxels = num_voxels.narrow(0, 0, size.size(0))
    out = (pos / size.view(1, -1)).to(torch.long)
    out = out * num_voxels.view(1, -1)
    return out.sum(1)


# torch_geometric.utils.scatter(src, index, dim=0, reduce='sum'|'mean') on CPU (= scatter_add_, ascending rows)
def scatter_sum(src, index, dim_size):
    idx = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
    return src.new_zeros((dim_size,) + tuple(src.shape[1:])).scatter_add_(0, idx, src)


def scatter_mean(src, index, dim_size=None):
    if dim_size is None:
        dim_size = int(index.max()) + 1 if index.numel() > 0 else 0
    count = src.new_zeros(dim_size).scatter_add_(0, index, src.new_ones(src.size(0)))
    count = count.clamp(min=1)
    out = scatter_sum(src, index, dim_size)
    return out / count.view(-1, *([1] * (src.dim() - 1)))


# torch_scatter.scatter_max(src, index, dim=0) -> values only (exact, order independent); empty rows stay 0
def scatter_max(src, index, dim_size=None):
    if dim_size is None:
        dim_size = int(index.max()) + 1 if index.numel() > 0 else 0
    idx = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
    out = src.new_zeros((dim_size,) + tuple(src.shape[1:]))
    return out.scatter_reduce_(0, idx, src, reduce="amax", include_self=False)


# torch_geometric BatchNorm -> nn.BatchNorm1d in eval mode (components.py:9-12)
def batch_norm_eval(x, weight, bias, running_mean, running_var, eps=1e-5):
    return torch.nn.functional.batch_norm(x, running_mean, running_var, weight, bias, False, 0.1, eps)
