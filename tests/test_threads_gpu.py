"""The reference runs under a threaded Flask server and holds no locks: one store / one encoder used
from several host threads on several CUDA streams must give the serial answers (the handles serialise their shared
workspaces on the device, _lib.HandleGuard)."""
import threading

import numpy as np
import pytest
import torch

from oracle import vit_oracle as vo

pytestmark = pytest.mark.gpu


def test_concurrent_search_and_encode(pkg, cuda):
    rng = np.random.default_rng(41)
    store = pkg.search.FrameStore.from_numpy(rng.standard_normal((200_000, 512)).astype(np.float32), device=cuda)
    visual = vo.build_visual("ViT-B/32", seed=1)
    enc = pkg.vit.VisionEncoder.from_state_dict(visual.state_dict(), "ViT-B/32", device=cuda, max_batch=16)
    queries = [torch.from_numpy(rng.standard_normal((q, 512)).astype(np.float32)).to(cuda) for q in (1, 7, 130, 33)]
    _, pix = vo.synthetic_frames(24, seed=5)
    batches = [pix[i:i + 6].to(cuda) for i in range(0, 24, 6)]
    serial_s = [store.search(q, 50) for q in queries]
    serial_e = [enc.encode(b, normalise=True) for b in batches]
    torch.cuda.synchronize()
    results, errors = {}, []

    def work(tid):
        try:
            stream = torch.cuda.Stream(device=cuda)
            with torch.cuda.stream(stream):
                for rep in range(6):
                    j = (tid + rep) % 4
                    s, i = store.search(queries[j], 50)
                    e = enc.encode(batches[j], normalise=True)
                    stream.synchronize()
                    results[(tid, rep)] = (j, s.clone(), i.clone(), e.clone())
        except Exception as exc:  # surfaced in the main thread
            errors.append(exc)

    threads = [threading.Thread(target=work, args=(t,)) for t in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    assert len(results) == 24
    for (tid, rep), (j, s, i, e) in results.items():
        assert torch.equal(s, serial_s[j][0]) and torch.equal(i, serial_s[j][1]), (tid, rep)
        assert torch.equal(e, serial_e[j]), (tid, rep)
