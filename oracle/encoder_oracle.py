"""TEST INFRASTRUCTURE (oracle): torch-CPU restatement of HierarchicalPoseEncoder.forward (models/network_utils.py:
119-178), the per-frame condition of the non-rigid MLP.  PINNED: tests/golden/make_pose_encoder_golden.py executes the
reference's OWN class (its source is lifted out of models/network_utils.py with `ast`, because the module itself imports
tinycudann and cannot be imported here) on seeded inputs and commits inputs, parameters and outputs as
tests/golden/ref_pose_encoder.npz; tests/test_oracle_encoder.py checks this restatement against them.
Only tests/ and __graft_entry__ import this."""
import torch
import torch.nn.functional as F

KTREE_PARENTS = [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 20, 21]  # network_utils.py:128-129


def encode(rots, Jtrs, W0, b0, Wa, ba, Wb, bb, rel_joints=False):
    """rots [24,9], Jtrs [24,3]; W0 [6,288], b0 [6]; Wa [24,19,19], ba [24,19]; Wb [24,6,19], bb [24,6] -> [144]."""
    J = Jtrs.clone()
    if rel_joints:  # :155-159
        J[1:] = Jtrs[1:] - Jtrs[KTREE_PARENTS[1:]]
        J = J.detach()  # formed under torch.no_grad() in the reference: no gradient reaches Jtrs in this mode
    glob = F.linear(torch.cat([rots.reshape(-1), J.reshape(-1)]), W0, b0)  # :161-162
    out = [None] * 24
    for j in range(24):
        p = KTREE_PARENTS[j]
        if p == -1:
            bone_l = J[j].norm().reshape(1)
            feat = glob
        else:
            bone_l = (J[j] if rel_joints else J[j] - J[p]).norm().reshape(1)
            feat = out[p]
        x = torch.cat([rots[j], J[j], bone_l, feat])
        out[j] = F.linear(F.relu(F.linear(x, Wa[j], ba[j])), Wb[j], bb[j])
    return torch.cat(out)
