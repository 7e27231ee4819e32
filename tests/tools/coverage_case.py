import sys
sys.path.insert(0, str(__import__("pathlib").Path(__file__).resolve().parent.parent.parent)); sys.path.insert(0, str(__import__("pathlib").Path(__file__).resolve().parent.parent.parent / "inference-arena_b200"))
import numpy as np, torch
from oracle import pipeline as O
from workloads import synth
from shared.processing import device as D
dev = "cuda"
# K1: copy / gather / linear / box2 / unaligned base / odd target
for (h, w, t) in [(480, 640, 640), (1080, 1920, 640), (333, 500, 640), (1280, 1280, 640), (123, 77, 333)]:
    f = synth.frames(2, h, w, seed=h)
    out = D.preprocess_batch_cuda(torch.from_numpy(f).to(dev), target=t)
    u8 = D.letterbox_batch_cuda(torch.from_numpy(f).to(dev), t)
    assert np.array_equal(out[0:1].cpu().numpy(), O.yolo_preprocess(f[0], target=t)[0])
raw = synth.frames(1, 251, 333, seed=9).reshape(-1)
buf = torch.from_numpy(raw).to(dev)
for off in (1, 13):
    D.preprocess_batch_cuda(buf[off:off + 250 * 333 * 3].view(1, 250, 333, 3))
# K2/K3: standard, dense global-sort path, ragged tiles, v8
for head in (synth.head_v5(2, seed=2), synth.head_v5_dense(1, seed=23, candidates=3000, anchors=3000, used_classes=2),
             synth.head_v5_dense(1, seed=31, candidates=40, anchors=129)):
    D.decode_nms_cuda(torch.from_numpy(head).to(dev), 0.25, 0.45, max_det=head.shape[1])
h8 = synth.head_v8_from_v5(synth.head_v5(1, seed=11, anchors=8400))
D.decode_nms_cuda(torch.from_numpy(h8).to(dev), 0.5, 0.45, layout="v8")
# K4: fan-out incl. edge ROIs on small and large frames; last-image / last-row ROIs (clamp path)
for (h, w) in [(480, 640), (1080, 1920), (375, 501)]:
    n = 3
    f = synth.frames(n, h, w, seed=3)
    dets, count = synth.detections(n, seed=31, kmax=8)
    dets[n - 1, 0, :4] = [0, 0, 640, 640]; dets[n - 1, 1, :4] = [600, 400, 640, 640]
    geom = D.letterbox_geometry(h, w, 640)
    crops, offs, rois = D.crop_batch_cuda(torch.from_numpy(f).to(dev), torch.from_numpy(dets).to(dev),
                                          torch.from_numpy(count).to(dev), geom)
torch.cuda.synchronize()
print("sanitizer case done")
