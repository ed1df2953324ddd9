"""Build-container-only helper: freeze the reference's size priors (tools/priors.pkl, a data file,
`priors_dims_per_cat` [50,2,3] fp64) as 3dod_b200/tdod/data/priors_dims_per_cat.npy so synthetic
frames can draw realistic (mean, std) rows on the GPU box, where /root/reference does not exist."""
import pickle, sys, types, os
import numpy as np

d2 = types.ModuleType('detectron2'); data = types.ModuleType('detectron2.data')
cat = types.ModuleType('detectron2.data.catalog')
class Metadata(types.SimpleNamespace):
    def __setstate__(self, s): self.__dict__.update(s)
cat.Metadata = Metadata; data.catalog = cat; d2.data = data
sys.modules.update({'detectron2': d2, 'detectron2.data': data, 'detectron2.data.catalog': cat})
p = pickle.load(open('/root/reference/tools/priors.pkl', 'rb'))
if isinstance(p, tuple):
    print([type(x) for x in p])
    p = [x for x in p if isinstance(x, dict)][0]
a = np.asarray(p['priors_dims_per_cat'], dtype=np.float64)
valid = ~np.all(a.reshape(a.shape[0], -1) == 1, axis=1)
print(a.shape, 'valid rows', int(valid.sum()))
out = os.path.join(os.path.dirname(__file__), '..', '3dod_b200', 'tdod', 'data')
os.makedirs(out, exist_ok=True)
np.save(os.path.join(out, 'priors_dims_per_cat.npy'), a)
