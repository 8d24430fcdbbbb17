"""GPU parity of the identity matching (utils/inference/video_processing.py:141-148) against the
reference's own torch lines (F.normalize, @, argmax(0), threshold) on the CPU."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import ghost_b200

pytestmark = pytest.mark.gpu


def reference_lines(face_embeds, target_embeds, th):
    target_embeds = F.normalize(target_embeds)                    # video_processing.py:126
    face_embeds = F.normalize(face_embeds)                        # :139
    similarity = face_embeds @ target_embeds.T                    # :141
    best_idxs = similarity.argmax(0).detach().cpu().numpy()       # :142
    matched = np.array([bool(similarity[b][i] > th) for i, b in enumerate(best_idxs)])   # :144
    return best_idxs, matched, similarity


@pytest.mark.parametrize("n_faces,n_targets", [(1, 1), (4, 4), (7, 3), (33, 5)])
def test_matches_reference_lines(n_faces, n_targets):
    g = torch.Generator().manual_seed(n_faces * 10 + n_targets)
    targets = torch.randn(n_targets, 512, generator=g) * 1.3
    faces = torch.randn(n_faces, 512, generator=g)
    for j in range(min(n_faces, n_targets)):                      # some faces are noisy copies of targets
        faces[(j * 3) % n_faces] = targets[j] * 0.7 + torch.randn(512, generator=g) * 0.4
    idx, ok, sim = ghost_b200.match_identities(faces.cuda(), targets.cuda(), 0.15)
    ridx, rok, rsim = reference_lines(faces, targets, 0.15)
    assert np.array_equal(idx, ridx) and np.array_equal(ok, rok)
    best = rsim[torch.from_numpy(ridx), torch.arange(n_targets)].numpy()
    assert np.abs(sim - best).max() <= 2e-6
    assert ok.any() and (n_targets == 1 or not ok.all() or n_faces >= n_targets)


def test_first_maximum_wins_and_zero_vectors():
    t = torch.zeros(2, 512)
    t[0, 0] = 2.0
    f = torch.zeros(5, 512)
    f[1, 0] = 1.0
    f[3, 0] = 3.0                                                # same direction as face 1: a tie after normalisation
    idx, ok, sim = ghost_b200.match_identities(f.cuda(), t.cuda(), 0.5)
    ridx, rok, _ = reference_lines(f, t, 0.5)
    assert np.array_equal(idx, ridx) and np.array_equal(ok, rok)   # zero target -> similarity 0 -> not matched
    assert idx[0] == 1 and ok[0] and not ok[1]
