"""The seven network variants of the reference (network/manifold*.prototxt) as a table, and an
emitter that writes an equivalent Caffe prototxt for each (the reference's files themselves are
not copied into this repo and do not exist on the GPU box)."""
from __future__ import annotations

# name -> (batch, input channels, concat groups or None, activation, fc outputs, TRAIN-phase tail)
#   groups name the Slice tops that are concatenated into conv1's input (rgb=0:3, dep=3:4, nor=4:7)
VARIANTS = {
    # network/manifold.prototxt:4,33-38,104
    "manifold": (330, 4, None, "ReLU", [256, 3], "triplet"),
    # network/manifold_wang.prototxt:4,35,104
    "manifold_wang": (300, 4, None, "ReLU", [256, 64], "none"),
    # network/manifold_rgb_16.prototxt:6-33,66,135
    "manifold_rgb_16": (300, 7, ["rgb"], "ELU", [256, 16], "none"),
    # network/manifold_rgbd_16.prototxt:20-33,66
    "manifold_rgbd_16": (300, 7, ["rgb", "dep"], "PReLU", [256, 16], "none"),
    # network/manifold_rgbnor_16.prototxt:20-33
    "manifold_rgbnor_16": (300, 7, ["rgb", "nor"], "ELU", [256, 16], "none"),
    # network/manifold_rgbdnor_16.prototxt:4,35,104
    "manifold_rgbdnor_16": (300, 7, None, "ELU", [256, 16], "none"),
    # network/manifold_rgbdnor_16_bin.prototxt:35,71,93,113-154
    "manifold_rgbdnor_16_bin": (300, 7, None, "PReLU", [256, 16, 512, 128], "binloss"),
}

_GROUP_CH = {"rgb": [0, 1, 2], "dep": [3], "nor": [4, 5, 6]}

# expected TEST-phase facts per variant: (selected channels, desc dim)
EXPECTED = {
    "manifold": ([0, 1, 2, 3], 3),
    "manifold_wang": ([0, 1, 2, 3], 64),
    "manifold_rgb_16": ([0, 1, 2], 16),
    "manifold_rgbd_16": ([0, 1, 2, 3], 16),
    "manifold_rgbnor_16": ([0, 1, 2, 4, 5, 6], 16),
    "manifold_rgbdnor_16": ([0, 1, 2, 3, 4, 5, 6], 16),
    "manifold_rgbdnor_16_bin": ([0, 1, 2, 3, 4, 5, 6], 128),
}


def _layer(name, ltype, bottoms, tops, body="", train_only=False):
    s = f'layer {{\n  name: "{name}"\n  type: "{ltype}"\n'
    for b in bottoms:
        s += f'  bottom: "{b}"\n'
    for t in tops:
        s += f'  top: "{t}"\n'
    s += body
    if train_only:
        s += "  include {phase: TRAIN}\n"
    return s + "}\n\n"


def prototxt(name: str) -> str:
    batch, c, groups, act, fcs, tail = VARIANTS[name]
    out = f'name: "MANIFOLD"\n\ninput: "data"\ninput_shape{{dim: {batch} dim: {c} dim: 64 dim: 64}}\n\n'
    src = "data"
    if groups is not None:
        out += _layer("slicer", "Slice", ["data"], ["rgb", "dep", "nor"],
                      "  slice_param {\n    axis: 1\n    slice_point: 3\n    slice_point: 4\n  }\n")
        out += _layer("concat", "Concat", groups, ["input"], "  concat_param { axis: 1 }\n")
        silenced = [g for g in ("rgb", "dep", "nor") if g not in groups]
        out += _layer("silencer", "Silence", silenced, [])
        src = "input"
    filler = '    weight_filler { type: "xavier" }\n    bias_filler { type: "constant" }\n'
    lr = "  param { lr_mult: 1 }\n  param { lr_mult: 2 }\n"
    out += _layer("conv1", "Convolution", [src], ["conv1"],
                  lr + "  convolution_param {\n    num_output: 16\n    kernel_size: 8\n" + filler + "  }\n")
    out += _layer("pool1", "Pooling", ["conv1"], ["pool1"],
                  "  pooling_param {\n    pool: MAX\n    kernel_size: 2\n    stride: 2\n  }\n")
    out += _layer("relu1", act, ["pool1"], ["pool1"])
    out += _layer("conv2", "Convolution", ["pool1"], ["conv2"],
                  lr + "  convolution_param {\n    num_output: 7\n    kernel_size: 5\n" + filler + "  }\n")
    out += _layer("pool2", "Pooling", ["conv2"], ["pool2"],
                  "  pooling_param {\n    pool: MAX\n    kernel_size: 2\n    stride: 2\n  }\n")
    out += _layer("relu2", act, ["pool2"], ["pool2"])
    fc_names = ["ip1", "ip2", "bin_ip1", "bin_ip2"]
    prev = "pool2"
    for i, nout in enumerate(fcs):
        fcn = fc_names[i]
        out += _layer(fcn, "InnerProduct", [prev], [fcn],
                      lr + f"  inner_product_param {{\n    num_output: {nout}\n" + filler + "  }\n")
        if i < len(fcs) - 1:
            out += _layer("relu_" + fcn, act, [fcn], [fcn])
        prev = fcn
    if tail == "triplet":   # TRAIN-only tail of network/manifold.prototxt:111-166
        out += _layer("slicer", "Slice", [prev], ["feat_triplets", "feat_pairs"],
                      "  slice_param {\n    axis: 0\n    slice_point: 198\n  }\n", train_only=True)
        out += _layer("tripletLoss", "TripletLoss", ["feat_triplets", "label_triplets"], ["loss1"],
                      "  triplet_loss_param{\n    margin: 0.1\n    losstype: 1\n    num_negatives: 1\n    use_pair: 0\n  }\n",
                      train_only=True)
    elif tail == "binloss":  # network/manifold_rgbdnor_16_bin.prototxt:156-163
        out += _layer("loss", "BinarizationLoss", [prev, prev], ["loss"], train_only=True)
    return out


def expected_blob_names(name: str):
    _, _, _, act, fcs, _ = VARIANTS[name]
    prelu = act == "PReLU"
    names = ["conv1.w", "conv1.b"] + (["relu1.a"] if prelu else [])
    names += ["conv2.w", "conv2.b"] + (["relu2.a"] if prelu else [])
    fc_names = ["ip1", "ip2", "bin_ip1", "bin_ip2"]
    for i in range(len(fcs)):
        names += [fc_names[i] + ".w", fc_names[i] + ".b"]
        if prelu and i < len(fcs) - 1:
            names.append("relu_" + fc_names[i] + ".a")
    return names
