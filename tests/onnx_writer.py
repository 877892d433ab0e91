"""Test aid: write ONNX protobufs (wire format by hand; the `onnx` package is not installable offline) that follow the
torch exporter's conventions for the two networks -- Linear weights renamed `onnx::MatMul_N` and stored [in, out] with
the bias on a separate Add, BatchNorms that follow a convolution folded into it (`onnx::Conv_N` + a new bias), the
remaining parameters under the ORIGINAL projects' module names (not ours), constants for rotary tables / GELU / FSQ
mixed in.  The graphs are structurally faithful (operator types, wiring of the weighted nodes), not executable."""
import struct

import numpy as np


def _varint(v: int) -> bytes:
    if v < 0:
        v += 1 << 64
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _key(f, wt):
    return _varint((f << 3) | wt)


def _ld(f, payload: bytes) -> bytes:
    return _key(f, 2) + _varint(len(payload)) + payload


def _s(f, text: str) -> bytes:
    return _ld(f, text.encode())


def tensor(name: str, arr: np.ndarray) -> bytes:
    arr = np.ascontiguousarray(arr)
    dt = {np.dtype(np.float32): 1, np.dtype(np.int64): 7}[arr.dtype]
    msg = b"".join(_key(1, 0) + _varint(d) for d in arr.shape)
    msg += _key(2, 0) + _varint(dt) + _s(8, name) + _ld(9, arr.tobytes())
    return msg


def _attr(name, val) -> bytes:
    msg = _s(1, name)
    if isinstance(val, float):
        msg += _key(2, 5) + struct.pack("<f", val) + _key(20, 0) + _varint(1)
    elif isinstance(val, int):
        msg += _key(3, 0) + _varint(val) + _key(20, 0) + _varint(2)
    elif isinstance(val, (list, tuple)):
        msg += _ld(8, b"".join(_varint(int(v)) for v in val)) + _key(20, 0) + _varint(7)
    elif isinstance(val, np.ndarray):
        msg += _ld(5, tensor("", val)) + _key(20, 0) + _varint(4)
    return msg


class Graph:
    def __init__(self):
        self.nodes, self.inits, self.n = [], [], 1000

    def fresh(self, prefix):
        self.n += 1
        return f"{prefix}_{self.n}"

    def init(self, name, arr):
        self.inits.append(tensor(name, np.asarray(arr)))
        return name

    def node(self, op, inputs, n_out=1, **attrs):
        outs = [self.fresh(f"/{op}_output") for _ in range(n_out)]
        msg = b"".join(_s(1, i) for i in inputs) + b"".join(_s(2, o) for o in outs)
        msg += _s(3, f"/{op}_{self.n}") + _s(4, op) + b"".join(_ld(5, _attr(k, v)) for k, v in attrs.items())
        self.nodes.append(msg)
        return outs[0]

    def save(self, path):
        g = b"".join(_ld(1, n) for n in self.nodes) + _s(2, "main_graph") + b"".join(_ld(5, t) for t in self.inits)
        model = _key(1, 0) + _varint(8) + _s(2, "pytorch") + _ld(7, g) + _ld(8, _s(1, "") + _key(2, 0) + _varint(17))
        with open(path, "wb") as f:
            f.write(model)


def _fold(w, bn, eps):
    mean, var, g, b = bn
    sc = g / np.sqrt(var + eps)
    return (w * sc.reshape(-1, *([1] * (w.ndim - 1)))).astype(np.float32), (b - mean * sc).astype(np.float32)


def write_campplus(path, st, cfg, fold=True, drop_layer=None, dense_as_gemm=False):
    """st: our synthetic state (synth_weights.campplus_state).  fold=True is what the exporter does in eval mode."""
    g, eps = Graph(), cfg.bn_eps
    bn_of = lambda n, affine=True: (st[n + ".running_mean"], st[n + ".running_var"],
                                     st[n + ".weight"] if affine else np.ones_like(st[n + ".running_mean"]),
                                     st[n + ".bias"] if affine else np.zeros_like(st[n + ".running_mean"]))

    def conv_bn(x, conv, bn, orig, affine=True, **attrs):
        w = st[conv + ".weight"]
        if fold:
            wf, bf = _fold(w, bn_of(bn, affine), eps)
            return g.node("Conv", [x, g.init(g.fresh("onnx::Conv"), wf), g.init(g.fresh("onnx::Conv"), bf)], **attrs)
        y = g.node("Conv", [x, g.init(orig + ".weight", w)], **attrs)
        return bn_node(y, bn, orig + "__bn", affine)       # unique per convolution

    def bn_node(x, bn, orig, affine=True):
        m, v, ga, be = bn_of(bn, affine)
        return g.node("BatchNormalization", [x, g.init(orig + ".weight", ga), g.init(orig + ".bias", be),
                                             g.init(orig + ".running_mean", m), g.init(orig + ".running_var", v)], epsilon=float(eps))

    x = g.node("Unsqueeze", [g.node("Transpose", ["input"], perm=[0, 2, 1]), g.init("axes0", np.array([1], np.int64))])
    x = g.node("Relu", [conv_bn(x, "head.conv1", "head.bn1", "head.conv1", kernel_shape=[3, 3], pads=[1, 1, 1, 1])])
    i = 0
    for li, nb in enumerate(cfg.fcm_blocks):
        for j in range(nb):
            pre, orig = f"head.blocks.{i}", f"head.layer{li + 1}.{j}"
            y = g.node("Relu", [conv_bn(x, pre + ".conv1", pre + ".bn1", orig + ".conv1", strides=[2 if j == 0 else 1, 1])])
            y = conv_bn(y, pre + ".conv2", pre + ".bn2", orig + ".conv2")
            sc = conv_bn(x, pre + ".sc_conv", pre + ".sc_bn", orig + ".shortcut.0", strides=[2, 1]) if j == 0 else x
            x = g.node("Relu", [g.node("Add", [y, sc])])
            i += 1
    x = g.node("Relu", [conv_bn(x, "head.conv2", "head.bn2", "head.conv2", strides=[2, 1])])
    x = g.node("Reshape", [x, g.init("shape0", np.array([0, -1, 0], np.int64))])
    x = g.node("Relu", [conv_bn(x, "tdnn", "tdnn_nl.0", "xvector.tdnn.linear", strides=[2], pads=[2, 2])])
    for b, nl in enumerate(cfg.block_layers):
        for l in range(nl):
            if drop_layer == (b, l):
                continue
            pre, orig = f"blocks.{b}.{l}", f"xvector.block{b + 1}.tdnnd{l + 1}"
            h = g.node("Relu", [bn_node(x, pre + ".nl1.0", orig + ".nonlinear1.batchnorm")])
            h = g.node("Relu", [conv_bn(h, pre + ".lin1", pre + ".nl2.0", orig + ".linear1")])
            y = g.node("Conv", [h, g.init(orig + ".cam_layer.linear_local.weight", st[pre + ".cam.local.weight"])],
                       dilations=[cfg.block_dilation[b]], pads=[cfg.block_dilation[b]] * 2)
            ctx = g.node("Add", [g.node("ReduceMean", [h], axes=[-1]), g.node("AveragePool", [h], kernel_shape=[cfg.seg_len], ceil_mode=1)])
            m = g.node("Relu", [g.node("Conv", [ctx, g.init(orig + ".cam_layer.linear1.weight", st[pre + ".cam.lin1.weight"]),
                                                g.init(orig + ".cam_layer.linear1.bias", st[pre + ".cam.lin1.bias"])])])
            m = g.node("Sigmoid", [g.node("Conv", [m, g.init(orig + ".cam_layer.linear2.weight", st[pre + ".cam.lin2.weight"]),
                                                   g.init(orig + ".cam_layer.linear2.bias", st[pre + ".cam.lin2.bias"])])])
            x = g.node("Concat", [x, g.node("Mul", [y, m])], axis=1)
        h = g.node("Relu", [bn_node(x, f"transits.{b}.nl.0", f"xvector.transit{b + 1}.nonlinear.batchnorm")])
        x = g.node("Conv", [h, g.init(f"xvector.transit{b + 1}.linear.weight", st[f"transits.{b}.lin.weight"])])
    x = g.node("Relu", [bn_node(x, "out_nl.0", "xvector.out_nonlinear.batchnorm")])
    mean = g.node("ReduceMean", [x], axes=[-1])
    std = g.node("Sqrt", [g.node("Mul", [g.node("ReduceMean", [g.node("Pow", [g.node("Sub", [x, mean]), g.init("two", np.array(2.0, np.float32))])], axes=[-1]),
                                        g.init("bessel", np.array(1.01, np.float32))])])
    x = g.node("Concat", [mean, std], axis=1)
    if dense_as_gemm:
        wf, bf = _fold(st["dense.weight"], bn_of("dense_bn", False), eps)
        x = g.node("Gemm", [g.node("Flatten", [x]), g.init(g.fresh("onnx::Gemm"), wf[:, :, 0]), g.init(g.fresh("onnx::Gemm"), bf)], transB=1)
    else:
        x = conv_bn(x, "dense", "dense_bn", "xvector.dense.linear", affine=False)
    g.save(path)


def write_s3tok(path, st, cfg, ln_eps=None, swap_kv=False, mlp_ln_eps=None):
    g, d = Graph(), cfg.d_model
    eps = float(cfg.ln_eps if ln_eps is None else ln_eps)
    eps2 = float(eps if mlp_ln_eps is None else mlp_ln_eps)

    def linear(x, name, orig, bias=True):
        y = g.node("MatMul", [x, g.init(g.fresh("onnx::MatMul"), np.ascontiguousarray(st[name + ".weight"].T))])
        return g.node("Add", [g.init(orig + ".bias", st[name + ".bias"]), y]) if bias else y

    def gelu(x):
        e = g.node("Erf", [g.node("Div", [x, g.init(g.fresh("c"), np.array(1.4142135, np.float32))])])
        return g.node("Mul", [g.node("Mul", [x, g.node("Add", [e, g.init(g.fresh("c"), np.array(1.0, np.float32))])]),
                              g.init(g.fresh("c"), np.array(0.5, np.float32))])

    x = gelu(g.node("Conv", ["feat", g.init("encoder.conv1.weight", st["conv1.weight"]), g.init("encoder.conv1.bias", st["conv1.bias"])],
                    strides=[2], pads=[1, 1]))
    x = gelu(g.node("Conv", [x, g.init("encoder.conv2.weight", st["conv2.weight"]), g.init("encoder.conv2.bias", st["conv2.bias"])],
                    strides=[2], pads=[1, 1]))
    x = g.node("Transpose", [x], perm=[0, 2, 1])
    cos = g.init("encoder.freqs_cos", np.ones((1, 64, 1, cfg.d_head), np.float32))
    sin = g.init("encoder.freqs_sin", np.zeros((1, 64, 1, cfg.d_head), np.float32))
    for i in range(cfg.n_layer):
        p, o = f"blocks.{i}.", f"encoder.blocks.{i}."
        h = g.node("LayerNormalization", [x, g.init(o + "attn_ln.weight", st[p + "attn_ln.weight"]),
                                          g.init(o + "attn_ln.bias", st[p + "attn_ln.bias"])], axis=-1, epsilon=eps)
        q = linear(h, p + "attn.query", o + "attn.query")
        if swap_kv:
            v = linear(h, p + "attn.value", o + "attn.value")
            k = linear(h, p + "attn.key", o + "attn.key", bias=False)
        else:
            k = linear(h, p + "attn.key", o + "attn.key", bias=False)
            v = linear(h, p + "attn.value", o + "attn.value")
        q = g.node("Add", [g.node("Mul", [q, cos]), g.node("Mul", [q, sin])])
        k = g.node("Add", [g.node("Mul", [k, cos]), g.node("Mul", [k, sin])])
        mem = g.node("Conv", [g.node("Transpose", [v], perm=[0, 2, 1]), g.init(o + "attn.fsmn_block.weight", st[p + "attn.fsmn.weight"])],
                     group=d, pads=[cfg.fsmn_kernel // 2] * 2)
        w = g.node("Softmax", [g.node("MatMul", [q, g.node("Transpose", [k], perm=[0, 1, 3, 2])])], axis=-1)
        a = linear(g.node("MatMul", [w, v]), p + "attn.out", o + "attn.out")
        x = g.node("Add", [x, g.node("Add", [a, mem])])
        h = g.node("LayerNormalization", [x, g.init(o + "mlp_ln.weight", st[p + "mlp_ln.weight"]),
                                          g.init(o + "mlp_ln.bias", st[p + "mlp_ln.bias"])], axis=-1, epsilon=eps2)
        x = g.node("Add", [x, linear(gelu(linear(h, p + "fc1", o + "mlp.0")), p + "fc2", o + "mlp.2")])
    h = g.node("Tanh", [linear(x, "project_down", "quantizer._codebook.project_down")])
    h = g.node("Round", [g.node("Mul", [h, g.init("fsq_scale", np.array(0.999, np.float32))])])
    g.node("ReduceSum", [g.node("Mul", [h, g.init("powers", (3.0 ** np.arange(8)).astype(np.float32))])])
    g.save(path)
