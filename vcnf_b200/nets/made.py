"""MADE conditioner (reference: normflow/nets/made.py:15-300): masked linear layers with
sequential degrees, residual blocks, output index = feature*multiplier + m."""
import torch
from torch import nn
from torch.nn import functional as F, init

from .. import ops
from ..utils import tile


def _get_input_degrees(in_features):
    return torch.arange(1, in_features + 1)


class MaskedLinear(nn.Linear):
    """Linear layer whose weight is multiplied by a fixed 0/1 degree mask (made.py:20-80)."""

    def __init__(self, in_degrees, out_features, autoregressive_features, random_mask, is_output,
                 bias=True, out_degrees_=None):
        super().__init__(in_features=len(in_degrees), out_features=out_features, bias=bias)
        mask, degrees = self._get_mask_and_degrees(in_degrees, out_features, autoregressive_features,
                                                   random_mask, is_output, out_degrees_)
        self.register_buffer("mask", mask)
        self.register_buffer("degrees", degrees)

    @classmethod
    def _get_mask_and_degrees(cls, in_degrees, out_features, autoregressive_features, random_mask,
                              is_output, out_degrees_=None):
        d = autoregressive_features
        if is_output:
            base = _get_input_degrees(d) if out_degrees_ is None else out_degrees_
            out_degrees = tile(base, out_features // d)
            return (out_degrees[..., None] > in_degrees).float(), out_degrees     # strict (made.py:60)
        if random_mask:
            lo = min(torch.min(in_degrees).item(), d - 1)
            out_degrees = torch.randint(low=lo, high=d, size=[out_features], dtype=torch.long)
        else:
            out_degrees = torch.arange(out_features) % max(1, d - 1) + min(1, d - 1)
        return (out_degrees[..., None] >= in_degrees).float(), out_degrees

    # W * mask held for the duration of a no-grad multi-pass call (MADE.frozen_masks): the D passes of the sampling
    # direction would otherwise recompute the same product D times per layer
    _masked = None

    def forward(self, x, relu_in=False, residual=None):
        """F.linear(act(x), W * mask, b) (+ residual) -- made.py:79-80.  fp32 CUDA operands run the tcgen05
        3xTF32 kernels (ReLU, bias and the residual add fused; hidden widths > 128 as contraction slices);
        fp64 models (`nfm.double()`) keep the library DGEMM: the tensor cores have no fp64-faithful path."""
        w = self._masked if self._masked is not None else self.weight * self.mask
        if ops.linear_supported(x, w):
            return ops.linear(x, w, self.bias, residual=residual, relu_in=relu_in)
        y = F.linear(F.relu(x) if relu_in else x, w, self.bias)
        return y if residual is None else residual + y


class MaskedFeedforwardBlock(nn.Module):
    def __init__(self, in_degrees, autoregressive_features, context_features=None, random_mask=False,
                 activation=F.relu, dropout_probability=0., use_batch_norm=False):
        super().__init__()
        features = len(in_degrees)
        self.batch_norm = nn.BatchNorm1d(features, eps=1e-3) if use_batch_norm else None
        if context_features is not None:
            raise NotImplementedError()
        self.linear = MaskedLinear(in_degrees, features, autoregressive_features, random_mask, False)
        self.degrees = self.linear.degrees
        self.activation = activation
        self.dropout = nn.Dropout(p=dropout_probability)

    def forward(self, inputs, context=None):
        if context is not None:
            raise NotImplementedError()
        h = self.batch_norm(inputs) if self.batch_norm else inputs
        return self.dropout(self.activation(self.linear(h)))


class MaskedResidualBlock(nn.Module):
    def __init__(self, in_degrees, autoregressive_features, context_features=None, random_mask=False,
                 activation=F.relu, dropout_probability=0., use_batch_norm=False, zero_initialization=True):
        if random_mask:
            raise ValueError("Masked residual block can't be used with random masks.")
        super().__init__()
        features = len(in_degrees)
        if context_features is not None:
            self.context_layer = nn.Linear(context_features, features)
        self.use_batch_norm = use_batch_norm
        if use_batch_norm:
            self.batch_norm_layers = nn.ModuleList([nn.BatchNorm1d(features, eps=1e-3) for _ in range(2)])
        first = MaskedLinear(in_degrees, features, autoregressive_features, False, False)
        second = MaskedLinear(first.degrees, features, autoregressive_features, False, False)
        self.linear_layers = nn.ModuleList([first, second])
        self.degrees = second.degrees
        if torch.all(self.degrees >= in_degrees).item() != 1:
            raise RuntimeError("In a masked residual block, the output degrees can't be"
                               " less than the corresponding input degrees.")
        self.activation = activation
        self.dropout = nn.Dropout(p=dropout_probability)
        if zero_initialization:
            init.uniform_(self.linear_layers[-1].weight, a=-1e-3, b=1e-3)
            init.uniform_(self.linear_layers[-1].bias, a=-1e-3, b=1e-3)

    def forward(self, inputs, context=None):
        is_relu = self.activation is F.relu or self.activation is torch.relu or isinstance(self.activation, nn.ReLU)
        if (context is None and not self.use_batch_norm and is_relu
                and (self.dropout.p == 0. or not self.training)):
            # both masked layers with the ReLU, the bias and the skip connection fused into the GEMM launches
            t = self.linear_layers[0](inputs, relu_in=True)
            return self.linear_layers[1](t, relu_in=True, residual=inputs)
        h = inputs
        if self.use_batch_norm:
            h = self.batch_norm_layers[0](h)
        h = self.linear_layers[0](self.activation(h))
        if self.use_batch_norm:
            h = self.batch_norm_layers[1](h)
        h = self.linear_layers[1](self.dropout(self.activation(h)))
        if context is not None:
            h = F.glu(torch.cat((h, self.context_layer(context)), dim=1), dim=1)
        return inputs + h


class MADE(nn.Module):
    def __init__(self, features, hidden_features, context_features=None, num_blocks=2, output_multiplier=1,
                 use_residual_blocks=True, random_mask=False, permute_mask=False, activation=F.relu,
                 dropout_probability=0., use_batch_norm=False, preprocessing=None):
        if use_residual_blocks and random_mask:
            raise ValueError("Residual blocks can't be used with random masks.")
        super().__init__()
        self.preprocessing = (lambda inputs: inputs) if preprocessing is None else preprocessing
        self._plain_inputs = preprocessing is None
        input_degrees_ = _get_input_degrees(features)
        if permute_mask:
            input_degrees_ = input_degrees_[torch.randperm(features)]
        self.initial_layer = MaskedLinear(input_degrees_, hidden_features, features, random_mask, False)
        if context_features is not None:
            self.context_layer = nn.Linear(context_features, hidden_features)
        block_cls = MaskedResidualBlock if use_residual_blocks else MaskedFeedforwardBlock
        blocks, prev = [], self.initial_layer.degrees
        for _ in range(num_blocks):
            blocks.append(block_cls(in_degrees=prev, autoregressive_features=features,
                                    context_features=context_features, random_mask=random_mask,
                                    activation=activation, dropout_probability=dropout_probability,
                                    use_batch_norm=use_batch_norm))
            prev = blocks[-1].degrees
        self.blocks = nn.ModuleList(blocks)
        self.final_layer = MaskedLinear(prev, features * output_multiplier, features, random_mask, True,
                                        out_degrees_=input_degrees_)

    def hidden(self, inputs, context=None):
        """Everything but the final masked layer (made.py:288-299 up to the last hidden activation)."""
        h = self.initial_layer(self.preprocessing(inputs))
        if context is not None:
            h = h + self.context_layer(context)
        for block in self.blocks:
            h = block(h, context)
        return h

    def forward(self, inputs, context=None):
        return self.final_layer(self.hidden(inputs, context))

    def frozen_masks(self):
        """Context manager: every masked layer computes W * mask once and reuses it until exit (no-grad callers
        whose weights do not change inside the block)."""
        net = self

        class _Frozen:
            def __enter__(self):
                with torch.no_grad():
                    for m in net.modules():
                        if isinstance(m, MaskedLinear):
                            m._masked = m.weight * m.mask

            def __exit__(self, *exc):
                for m in net.modules():
                    if isinstance(m, MaskedLinear):
                        m._masked = None
                return False
        return _Frozen()

    def sampling_kernel_ok(self):
        """vcnf_made_inverse covers: residual blocks (<= 2) with ReLU, no preprocessing / context / batch norm / active
        dropout, fp32 CUDA parameters, widths that split over 8 warps, <= 8 hidden units per degree, one degree
        vector shared by all hidden layers."""
        if not getattr(self, "_plain_inputs", False) or hasattr(self, "context_layer"):
            return False
        if not all(isinstance(b, MaskedResidualBlock) for b in self.blocks) or not 1 <= 1 + 2 * len(self.blocks) <= 5:
            return False
        w = self.initial_layer.weight
        if not w.is_cuda or w.dtype != torch.float32:
            return False
        H, D = w.shape
        P = self.final_layer.weight.shape[0] // D
        nl, Pp = 1 + 2 * len(self.blocks), (P + 3) // 4 * 4
        smem = ((D + (nl - 1) * H) * 8 + nl * 8 + (H + 1) * Pp + (2 * D + nl * H + 16 * 32 + 32) * 16) * 4
        if D % 16 or H % 16 or P > 32 or smem > 220 * 1024:
            return False
        for b in self.blocks:
            relu = b.activation is F.relu or b.activation is torch.relu or isinstance(b.activation, nn.ReLU)
            if b.use_batch_norm or not relu or (b.dropout.p > 0. and b.training) or hasattr(b, "context_layer"):
                return False
        return True

    def pack_for_sampling(self):
        """Operands of vcnf_made_inverse (include/vcnf_b200.h): per-degree bands of the masked weights of every hidden
        layer, the final layer's rows grouped by feature in degree order.  Built per call (the weights are frozen only
        for its duration); None when the degree structure does not fit the kernel."""
        dev = self.initial_layer.weight.device
        H, D = self.initial_layer.weight.shape
        fl = self.final_layer
        P = fl.weight.shape[0] // D
        hdeg = self.initial_layer.degrees
        layers = [self.initial_layer]
        for b in self.blocks:
            layers += list(b.linear_layers)
        cache = self.__dict__.setdefault("_pack_index", {})
        key = (str(dev), hdeg.data_ptr(), hdeg._version, fl.degrees.data_ptr(), fl.degrees._version)
        if cache.get("key") != key:
            if any(not torch.equal(l.degrees, hdeg) for l in layers[1:]):
                cache.update(key=key, ok=False)
            else:
                hd = hdeg.cpu().tolist()
                bands = [[j for j, d in enumerate(hd) if d == t] for t in range(D)]
                fdeg = fl.degrees[::P].cpu()
                ok = max(len(b) for b in bands) <= 8 and max(hd) <= D - 1 and sorted(fdeg.tolist()) == list(range(1, D + 1))
                idx = torch.full((D, 8), -1, dtype=torch.int32)
                for t, b in enumerate(bands):
                    if b and len(b) <= 8:
                        idx[t, :len(b)] = torch.tensor(b, dtype=torch.int32)
                order = torch.argsort(fdeg, stable=True).to(torch.int32)
                cache.update(key=key, ok=ok, band_idx=idx.to(dev), band_units=max(len(b) for b in bands), gather=torch.where(idx < 0, H, idx).long().view(-1).to(dev),
                             order=order.to(dev), order_l=order.long().to(dev))
        if not cache["ok"]:
            return None
        gather = cache["gather"]
        Pp = (P + 3) // 4 * 4
        Wb, bb = [], []
        with torch.no_grad():
            for l in layers:
                wm = l._masked if l._masked is not None else l.weight * l.mask
                hin = wm.shape[1]
                ext = torch.cat([wm, wm.new_zeros(1, hin)])
                Wb.append(ext[gather].view(D, 8, hin).permute(0, 2, 1).reshape(D, hin * 8))
                bias = l.bias if l.bias is not None else wm.new_zeros(H)
                bb.append(torch.cat([bias, bias.new_zeros(1)])[gather].view(D, 8))
            band = torch.cat(Wb + bb, dim=1).contiguous()              # one contiguous block per pass
            wf = fl._masked if fl._masked is not None else fl.weight * fl.mask
            wf = wf.view(D, P, H)[cache["order_l"]]
            Wf = F.pad(wf, (0, 0, 0, Pp - P)).permute(0, 2, 1).reshape(D, H * Pp)
            bfv = fl.bias if fl.bias is not None else wf.new_zeros(D * P)
            bf = F.pad(bfv.view(D, P)[cache["order_l"]], (0, Pp - P))
            fin = torch.cat([Wf, bf], dim=1).contiguous()
        return dict(H=H, P=P, n_layers=len(layers), band=band, fin=fin, band_idx=cache["band_idx"], order=cache["order"],
                    band_units=cache["band_units"])

    def final_rows(self, h, lo, hi):
        """Outputs [lo, hi) of the final layer only -- the same dot products as `forward(...)[:, lo:hi]`."""
        fl = self.final_layer
        w = (fl._masked if fl._masked is not None else fl.weight * fl.mask)[lo:hi]
        b = fl.bias[lo:hi] if fl.bias is not None else None
        if ops.linear_supported(h, w):
            return ops.linear(h, w, b, pad_out=True)
        return F.linear(h, w, b)
