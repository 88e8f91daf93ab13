"""The caller of the hot path: a per-sequence evaluation loop with the shape of the reference's
``Evaluator.evaluating`` (networks/managers/evaluator.py:276-568, single scale, no flip), its input-size rule, its
sequence work queue and its palette PNG output -- so that the engines can be driven the way ``tools/eval.py`` drives
the reference's, over any iterable of sequences.

What is restated here and from where:
  * `restrict_size`      MultiRestrictSize (dataloaders/video_transforms.py:573-615, align_corners = True: every side
                         becomes 16 k + 1, long edge capped at max_long_edge first);
  * `memory_gap`         gap = max(round(n / 30), 5) (evaluator.py:330-335, NO_MEMORY_GAP divides it by 4);
  * `evaluate_sequence`  frame 0 -> add_reference_frame with the label resized (nearest) to the image; frames > 0 ->
                         match_propogate_one_frame(output_size = original size), softmax, argmax, new objects of a
                         frame's own label merged in and re-referenced (:485-509), else update_memory with the
                         label resized (nearest) to the engine's input size (:510-523);
  * `SequenceQueue`      rank 0 puts every sequence index, then one 'END' per worker; workers take the next index
                         when they reach theirs (:276-295) -- here over torch.multiprocessing queues or, in one process,
                         a plain deque; `shard.assign_sequences` is the static equivalent used by bench.py;
  * `palette` / `save_mask`  the DAVIS palette PNGs of utils/image.py:5-57, 90-106.
"""
import collections
import os
import threading

import numpy as np
import torch
import torch.nn.functional as F


# ------------------------------------------------------------------------------------------------ sizes and gaps
def restrict_size(h, w, max_long_edge=800 * 1.3, max_stride=16, align_corners=True, scale=1.0, min_short_edge=None):
    """Input size the evaluator feeds the model for an (h, w) frame (video_transforms.py:573-615; the default long-edge
    cap is tools/eval.py:78,128: 480 * 1.3 * 800 / 480 = 1040).  `min_short_edge` selects the short-edge rule instead
    (TEST_MIN_SIZE; the reference asserts that only one of the two is set)."""
    sc = None
    if min_short_edge is not None:
        short_edge = w if h > w else h
        if short_edge > min_short_edge:
            sc = float(min_short_edge) / short_edge
    elif max_long_edge is not None:
        long_edge = h if h > w else w
        if long_edge > max_long_edge:
            sc = float(max_long_edge) / long_edge
    new_h, new_w = (h, w) if sc is None else (sc * h, sc * w)
    new_h, new_w = int(new_h * scale), int(new_w * scale)
    if align_corners:
        if (new_h - 1) % max_stride != 0:
            new_h = int(np.around((new_h - 1) / max_stride) * max_stride + 1)
        if (new_w - 1) % max_stride != 0:
            new_w = int(np.around((new_w - 1) / max_stride) * max_stride + 1)
    else:
        if new_h % max_stride != 0:
            new_h = int(np.around(new_h / max_stride) * max_stride)
        if new_w % max_stride != 0:
            new_w = int(np.around(new_w / max_stride) * max_stride)
    return new_h, new_w


def memory_gap(num_frames, no_memory_gap=False):
    """evaluator.py:330-335."""
    gap = max(int(round(num_frames / 30)), 5)
    if no_memory_gap:
        gap = int(round(gap / 4))
    return gap


# ------------------------------------------------------------------------------------------------ palette PNGs
def palette():
    """The 256-entry palette of utils/image.py:5-57 as a flat list of 768 ints: the DAVIS colours for the first 22
    labels (the PASCAL-VOC bit-interleaving construction with 192 written as 191), grey (i, i, i) from 22 on."""
    pal = []
    for i in range(256):
        if i >= 22:
            pal += [i, i, i]
            continue
        c, rgb = i, [0, 0, 0]
        for j in range(8):
            for ch in range(3):
                rgb[ch] |= ((c >> ch) & 1) << (7 - j)
            c >>= 3
        pal += [191 if v == 192 else v for v in rgb]
    return pal


def unsqueeze_labels(mask, squeeze_idx):
    """Map the engine's dense object numbers 1..n back to the dataset's label values (utils/image.py:91-97)."""
    if squeeze_idx is None:
        return mask
    out = mask * 0
    for idx in range(1, len(squeeze_idx)):
        out += ((mask == idx) * squeeze_idx[idx]).astype(np.uint8)
    return out


def write_mask_png(mask, path, squeeze_idx=None):
    """utils/image.py:90-101 (_save_mask): uint8 label map [H, W] -> palette PNG."""
    from PIL import Image
    mask = unsqueeze_labels(np.asarray(mask, dtype=np.uint8), squeeze_idx)
    im = Image.fromarray(mask).convert("P")
    im.putpalette(palette())
    im.save(path)


def save_mask(mask_tensor, path, squeeze_idx=None, threads=None):
    """utils/image.py:103-106: the write happens on a thread of its own (collected in ``threads`` if given)."""
    mask = mask_tensor.cpu().numpy().astype("uint8")
    t = threading.Thread(target=write_mask_png, args=[mask, path, squeeze_idx])
    t.start()
    if threads is not None:
        threads.append(t)
    return t


# ------------------------------------------------------------------------------------------------ work queue
class SequenceQueue:
    """The evaluator's sequence hand-out (evaluator.py:276-295): one producer (rank 0) puts every index followed by
    one 'END' per worker; each worker holds the next index it has to process and skips the others."""

    def __init__(self, n_workers=1, mp_queue=None):
        self.n_workers = n_workers
        self.q = mp_queue if mp_queue is not None else collections.deque()

    def _put(self, x):
        self.q.put(x) if hasattr(self.q, "put") else self.q.append(x)

    def _get(self):
        return self.q.get() if hasattr(self.q, "get") else self.q.popleft()

    def fill(self, n_sequences):
        for i in range(n_sequences):
            self._put(i)
        for _ in range(self.n_workers):
            self._put("END")

    def worker(self, n_sequences):
        """Yields the sequence indices this worker processes."""
        coming = self._get()
        for seq_idx in range(n_sequences):
            if coming == "END":
                return
            if coming != seq_idx:
                continue
            coming = self._get()
            yield seq_idx


# ------------------------------------------------------------------------------------------------ per-sequence loop
def evaluate_sequence(engine, frames, first_label, obj_num, original_size=None, later_labels=None,
                      result_dir=None, names=None, obj_idx=None, no_memory_gap=False, prefetch=True, on_mask=None):
    """One video through the engine the way Evaluator.evaluating does it (evaluator.py:318-568).

    frames        sequence of image tensors [1, 3, h, w] (already at the model's input size, see `restrict_size`)
    first_label   [1, 1, H, W] label map of frame 0 (any size: resized nearest to the image, :393-396)
    obj_num       number of objects in frame 0
    original_size (H0, W0) the masks are produced at (default: the image size)
    later_labels  {frame index: [1, 1, H0, W0] label map}: new objects appearing later (:485-509)
    result_dir    write `names[f] + '.png'` palette PNGs there (utils/image.py:90-106)
    Returns the list of uint8 label maps of frames 1..n-1 (CPU) unless `on_mask` consumes them."""
    n = len(frames)
    dev = engine.aot_engines[0].device if getattr(engine, "aot_engines", None) else None
    engine.restart_engine()
    engine.long_term_mem_gap = memory_gap(n, no_memory_gap)
    img0 = frames[0] if dev is None else frames[0].to(dev, non_blocking=True)
    if dev is None:
        dev = img0.device
    hw = tuple(img0.shape[-3:-1]) if img0.dtype == torch.uint8 else tuple(img0.shape[2:])
    out_size = tuple(original_size) if original_size is not None else hw
    lab0 = F.interpolate(first_label.to(dev).float(), size=hw, mode="nearest").int()
    engine.add_reference_frame(img0, lab0, frame_step=0, obj_nums=[obj_num])
    masks, threads = [], []
    obj_nums = [obj_num]
    if result_dir is not None:
        os.makedirs(result_dir, exist_ok=True)
    single = len(engine.aot_engines) == 1
    nxt = frames[1].to(dev, non_blocking=True) if n > 1 else None
    for f in range(1, n):
        cur = nxt
        nxt = frames[f + 1].to(dev, non_blocking=True) if f + 1 < n else None
        kw = {"next_img": nxt} if (prefetch and single and nxt is not None) else {}
        logit = engine.match_propogate_one_frame(cur, output_size=out_size, **kw)
        prob = torch.softmax(logit, dim=1)
        label = torch.argmax(prob, dim=1, keepdim=True).float()
        new_label = None if later_labels is None else later_labels.get(f)
        if new_label is not None:
            new_label = new_label.to(dev).float()
            keep = (new_label == 0).float()
            label = label * keep + new_label * (1 - keep)
            obj_nums = [int(label.max().item())]
            engine.add_reference_frame(cur, F.interpolate(label, size=engine.input_size_2d, mode="nearest"),
                                       obj_nums=obj_nums, frame_step=f)
        else:
            engine.update_memory(F.interpolate(label, size=engine.input_size_2d, mode="nearest"))
        m = label[0, 0].to(torch.uint8)
        if result_dir is not None:
            name = names[f] if names is not None else "%05d" % f
            save_mask(m, os.path.join(result_dir, name + ".png"), obj_idx, threads)
        if on_mask is not None:
            on_mask(f, m)
        else:
            masks.append(m.cpu())
    for t in threads:
        t.join()
    return masks


def evaluate(engine, sequences, queue=None, **kw):
    """Run `evaluate_sequence` over an indexable collection of sequence dicts (keys = its arguments), taking the
    indices from a `SequenceQueue` when one is given (multi-process evaluation) and all of them otherwise."""
    order = queue.worker(len(sequences)) if queue is not None else range(len(sequences))
    results = {}
    for i in order:
        results[i] = evaluate_sequence(engine, **sequences[i], **kw)
    return results
