import numpy as np, pose_estimation_b200 as pb
from pose_estimation_b200 import synth
g=np.load('tests/golden/callsites_ref.npz')
coor, uv, Ks = g["crb_coor"], g["crb_uv"], g["crb_K"]
wrap = pb.choose_ransac_batch(coor, uv, Ks, n=400)
out = pb.choose_ransac_batch(coor, uv, Ks, n=125)
for j in range(4):
    sp = synth.make_frame_sparse(0x5EED, 203, j, **(dict(noise_m=0.0, outlier_min_px=5.0) if j < 2 else {}))
    truth=np.nonzero(sp['inlier'][:300])[0]
    cvi=np.unique(g[f'crb_wrap{j}']); ours=np.unique(wrap[j])
    print(j,'truth',truth.size,'cv2',cvi.size,'ours',ours.size,'cv2 in ours',np.isin(cvi,ours).mean(),'ours in truth',np.isin(ours,truth).mean(), 'sel in ours', np.isin(out[j,0],ours).mean())
