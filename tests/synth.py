"""Seeded synthetic inputs of SURVEY.md 8(d), shared by the tests, smoke() and bench.py.
numpy only (no torch, no oracle) so that every consumer sees bit-identical inputs."""
import numpy as np

FPN_SCALES = (0.25, 0.125, 0.0625, 0.03125)
IMG_W, IMG_H = 1333, 800
PAD_W, PAD_H = 1344, 800  # SIZE_DIVISIBILITY 32


def fpn_shapes(n_images, channels=256, pad_w=PAD_W, pad_h=PAD_H):
    return [(n_images, channels, int(round(pad_h * s)), int(round(pad_w * s))) for s in FPN_SCALES]


def fpn_features(n_images, channels=256, seed=1234, pad_w=PAD_W, pad_h=PAD_H):
    out = []
    for lvl, shp in enumerate(fpn_shapes(n_images, channels, pad_w, pad_h)):
        rng = np.random.default_rng(seed + lvl)
        out.append(rng.standard_normal(shp, dtype=np.float32))
    return out


ADVERSARIAL_BOXES = np.array([
    [50.0, 50.0, 50.0, 50.0],        # zero size
    [5.0, 5.0, 6.0, 6.0],            # 1 px
    [0.0, 0.0, IMG_W - 1, IMG_H - 1],  # full image
    [1200.0, 700.0, 1332.0, 799.0],  # hugging the corner
], np.float32)


def proposals(n, seed, img_w=IMG_W, img_h=IMG_H, smin=16.0, smax=800.0, adversarial=True):
    """sqrt(area) log-uniform in [smin,smax], aspect in {1/2,1,2}*U(0.8,1.25), centres uniform,
    clipped to [0,W-1]x[0,H-1] (bounding_box.py:270-273)."""
    rng = np.random.default_rng(seed)
    s = np.exp(rng.uniform(np.log(smin), np.log(smax), n))
    ar = rng.choice([0.5, 1.0, 2.0], n) * rng.uniform(0.8, 1.25, n)
    w = s * np.sqrt(ar)
    h = s / np.sqrt(ar)
    cx = rng.uniform(0, img_w, n)
    cy = rng.uniform(0, img_h, n)
    b = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2], 1)
    b[:, 0::2] = np.clip(b[:, 0::2], 0, img_w - 1)
    b[:, 1::2] = np.clip(b[:, 1::2], 0, img_h - 1)
    b = b.astype(np.float32)
    if adversarial and n >= 8:
        b[-4:] = ADVERSARIAL_BOXES
    return b


def rois_for_images(n_images, per_image, seed=4321, **kw):
    """[n_images*per_image, 5] = (image, x1,y1,x2,y2), image-major."""
    rows = []
    for i in range(n_images):
        b = proposals(per_image, seed + i, **kw)
        rows.append(np.concatenate([np.full((per_image, 1), i, np.float32), b], 1))
    return np.concatenate(rows, 0).astype(np.float32)


def nms_boxes(n, seed, img_w=IMG_W, img_h=IMG_H, clusters=256, tie_heavy=False):
    """Clustered boxes (realistic 70-95% suppression) and scores (SURVEY.md 8(d) cfg 2)."""
    rng = np.random.default_rng(seed)
    centres = rng.uniform(0, 1, (clusters, 2)) * [img_w, img_h]
    which = rng.integers(0, clusters, n)
    c = centres[which] + rng.normal(0, 12.0, (n, 2))
    # sizes are a property of the cluster (objects), jittered per box
    cs = np.exp(rng.uniform(np.log(16), np.log(512), clusters))
    car = rng.choice([0.5, 1.0, 2.0], clusters)
    s = cs[which] * rng.uniform(0.8, 1.25, n)
    ar = car[which] * rng.uniform(0.8, 1.25, n)
    w = s * np.sqrt(ar)
    h = s / np.sqrt(ar)
    b = np.stack([c[:, 0] - w / 2, c[:, 1] - h / 2, c[:, 0] + w / 2, c[:, 1] + h / 2], 1)
    b[:, 0::2] = np.clip(b[:, 0::2], 0, img_w - 1)
    b[:, 1::2] = np.clip(b[:, 1::2], 0, img_h - 1)
    scores = rng.permutation(np.linspace(0.01, 0.99, n)).astype(np.float32)
    if tie_heavy:
        scores = (np.floor(scores * 256) / 256).astype(np.float32)
    labels = rng.integers(1, 81, n).astype(np.float32)
    return b.astype(np.float32), scores, labels


def yolact_priors(max_size=550):
    """Prior boxes exactly as Yolact.jittor/yolact.py:198-238 for yolact_resnet50
    (conv sizes 69,35,18,9,5; scales 24..384; aspect ratios [1,1/2,2]; square anchors)."""
    out = []
    for conv, scale in zip((69, 35, 18, 9, 5), (24, 48, 96, 192, 384)):
        for j in range(conv):
            for i in range(conv):
                x = (i + 0.5) / conv
                y = (j + 0.5) / conv
                for ar in (1, 0.5, 2):
                    ar = np.sqrt(ar)
                    w = scale * ar / max_size
                    h = w  # use_square_anchors
                    out.append([x, y, w, h])
    return np.asarray(out, np.float32)


def yolact_image(seed, priors, num_classes=81, mask_dim=32, proto_hw=138, hot_frac=0.02):
    """One image of Yolact head outputs (SURVEY.md 8(d) cfg 4)."""
    rng = np.random.default_rng(seed)
    P = priors.shape[0]
    loc = rng.normal(0, 0.5, (P, 4)).astype(np.float32)
    logits = rng.standard_normal((P, num_classes)).astype(np.float32)
    logits[:, 0] += 6.0
    hot = rng.random(P) < hot_frac
    cls = rng.integers(1, num_classes, P)
    logits[hot, cls[hot]] += 8.0
    e = np.exp(logits - logits.max(1, keepdims=True))
    conf = (e / e.sum(1, keepdims=True)).astype(np.float32)
    mask = np.tanh(rng.standard_normal((P, mask_dim))).astype(np.float32)
    proto = np.maximum(rng.standard_normal((proto_hw, proto_hw, mask_dim)), 0).astype(np.float32)
    return loc, conf, mask, proto


def paste_inputs(n, seed, M=28, img_w=IMG_W, img_h=IMG_H):
    """Mask probabilities (smoothed) and fractional boxes (SURVEY.md 8(d) cfg 5)."""
    rng = np.random.default_rng(seed)
    raw = rng.normal(0, 2.0, (n, M + 2, M + 2))
    sm = sum(raw[:, dy:dy + M, dx:dx + M] for dy in range(3) for dx in range(3)) / 9.0
    masks = (1.0 / (1.0 + np.exp(-sm))).astype(np.float32)
    b = proposals(n, seed + 17, img_w, img_h, smin=24.0, smax=600.0, adversarial=False)
    b = b + rng.uniform(0, 1, b.shape).astype(np.float32) * 0.9
    if n >= 8:
        b[-4:] = np.array([[-30.5, -20.25, 100.75, 90.5], [700.2, 400.1, 700.9, 400.6],
                           [0, 0, img_w - 1, img_h - 1], [1300.5, 780.5, 1400.0, 850.0]], np.float32)
    return masks, b.astype(np.float32)


def rpn_level(n_images, A, H, W, stride, seed, size=None, ties=False, img_w=IMG_W, img_h=IMG_H):
    """One FPN level of RPN head outputs + its anchors.
    anchors [n_images, H*W*A, 4] in the reference's (h, w, a) order: A aspect ratios
    (0.5, 1, 2, ...) of a `size`-pixel box (legacy +1/-1 corners) centred on each stride cell;
    objectness [N,A,H,W] logits (quantised to 1/4 when ties=True: many equal scores);
    box_regression [N,4A,H,W] with a few dw/dh above bbox_xform_clip and a few collapsing boxes."""
    rng = np.random.default_rng(seed)
    size = float(size if size is not None else stride * 8)
    ratios = np.array([0.5, 1.0, 2.0, 0.25, 4.0][:A] if A <= 5 else np.linspace(0.25, 4, A))
    ws = np.round(np.sqrt(size * size / ratios))
    hs = np.round(ws * ratios)
    cell = np.stack([-(ws - 1) / 2, -(hs - 1) / 2, (ws - 1) / 2, (hs - 1) / 2], 1) + (stride - 1) / 2
    sy, sx = np.meshgrid(np.arange(H) * stride, np.arange(W) * stride, indexing="ij")
    shifts = np.stack([sx, sy, sx, sy], -1).reshape(-1, 1, 4)
    anchors = (shifts + cell[None]).reshape(-1, 4).astype(np.float32)
    anchors = np.broadcast_to(anchors, (n_images,) + anchors.shape).copy()
    obj = rng.standard_normal((n_images, A, H, W)).astype(np.float32) * 2.0 - 1.0
    if ties:
        obj = np.round(obj * 4.0) / 4.0
    reg = (rng.standard_normal((n_images, 4 * A, H, W)) * 0.4).astype(np.float32)
    flat = reg.reshape(-1)
    hot = rng.choice(flat.size, max(4, flat.size // 200), replace=False)
    flat[hot] = rng.choice([5.0, 4.2, -6.0, -9.0], hot.size).astype(np.float32)
    im_wh = np.tile(np.array([[img_w, img_h]], np.float32), (n_images, 1))
    if n_images > 1:
        im_wh[1] = (int(img_w * 0.75), int(img_h * 0.9))  # images of a batch differ in size
    return anchors, obj.astype(np.float32), reg, im_wh


def box_head_batch(n_images, per_image, num_classes=81, seed=0, ties=False, objects=12,
                   img_w=IMG_W, img_h=IMG_H):
    """Box-head outputs for a batch: per image `objects` ground-truth-like
    boxes with a class each; 70 % of the proposals are jittered copies of them (high logit on the
    object's class -> real per-class suppression), the rest are background.  ties=True draws the
    foreground logit rows from a pool of 16 rows: many EQUAL scores on different boxes, which
    exercises the sort's tie order and the kthvalue cut.
    Returns logits [R,C], regression [R,4C], proposals [R,4], per-image counts, im_wh [N,2]."""
    rng = np.random.default_rng(seed)
    C = num_classes
    counts = [per_image if i % 2 == 0 else max(1, per_image - 37) for i in range(n_images)]
    pool = rng.standard_normal((16, C)).astype(np.float32)
    pool[:, 0] = rng.uniform(0, 2, 16)
    logits, props = [], []
    for b, n in enumerate(counts):
        obj = proposals(objects, seed * 131 + b, img_w, img_h, smin=40.0, smax=500.0, adversarial=False)
        cls = rng.integers(1, C, objects)
        lg = rng.standard_normal((n, C)).astype(np.float32)
        lg[:, 0] += 6.0
        pr = proposals(n, seed * 977 + b, img_w, img_h, adversarial=False)
        fg = rng.random(n) < 0.7
        which = rng.integers(0, objects, n)
        for i in np.nonzero(fg)[0]:
            o = obj[which[i]]
            w, h = o[2] - o[0], o[3] - o[1]
            pr[i] = o + rng.normal(0, 0.08, 4) * np.array([w, h, w, h])
            if ties:
                lg[i] = pool[rng.integers(0, 16)]
                lg[i, cls[which[i]]] = 4.0
            else:
                lg[i, 0] = rng.uniform(0, 3)
                lg[i, cls[which[i]]] = rng.uniform(2, 6)
                lg[i, rng.integers(1, C)] += rng.uniform(0, 4)  # a second plausible class
        pr[:, 0::2] = np.clip(np.sort(pr[:, 0::2], 1), 0, img_w - 1)
        pr[:, 1::2] = np.clip(np.sort(pr[:, 1::2], 1), 0, img_h - 1)
        logits.append(lg)
        props.append(pr.astype(np.float32))
    logits = np.concatenate(logits).astype(np.float32)
    props = np.concatenate(props).astype(np.float32)
    reg = (rng.standard_normal((logits.shape[0], 4 * C)) * 1.5).astype(np.float32)
    reg[::97, 2::4] = 30.0  # dw above the clamp
    im_wh = np.tile(np.array([[img_w, img_h]], np.float32), (n_images, 1))
    if n_images > 1:
        im_wh[1] = (int(img_w * 0.75), int(img_h * 0.9))
    return logits, reg, props, counts, im_wh


def embed_mask_inputs(n_center, h, w, dim, seed=0):
    """EmbedMask mask_prob inputs: pixel embeddings [h*w, dim], centre
    embeddings, sigmas, int boxes (end exclusive, inside the map) and their areas."""
    rng = np.random.default_rng(seed)
    ep = rng.standard_normal((h * w, dim)).astype(np.float32)
    ec = rng.standard_normal((n_center, dim)).astype(np.float32)
    sg = rng.uniform(0.05, 2.0, n_center).astype(np.float32)
    x1 = rng.integers(0, w - 2, n_center)
    y1 = rng.integers(0, h - 2, n_center)
    x2 = np.minimum(x1 + rng.integers(1, w, n_center), w)
    y2 = np.minimum(y1 + rng.integers(1, h, n_center), h)
    bx = np.stack([x1, y1, x2, y2], 1).astype(np.int32)
    if n_center > 2:
        bx[-1] = (0, 0, w, h)
    areas = ((bx[:, 2] - bx[:, 0]) * (bx[:, 3] - bx[:, 1])).astype(np.int32)
    return ep, ec, sg, bx, areas, w
