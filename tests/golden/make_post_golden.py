#!/usr/bin/env python
"""Golden vectors for the post-processing rows (f2), from the reference's own code.

    python tests/golden/make_post_golden.py          (build container only)

* lifter: the function ``lifter`` of /root/reference/software/lift.py, taken from the
  file's source and executed (the module itself imports matplotlib, which is absent, so
  only that function definition is compiled), then ``astype(np.int16)`` as at :39;
* standardisation: ``sklearn.preprocessing.scale(spec, axis=1)`` and the int16 cast of
  software/genlibrosa.py:25-28, run on the transposed matrix.
Input: the integer cepstra of the bundled clip (first 400 frames), from the C oracle.
"""
import ast
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def reference_lifter():
    src = open("/root/reference/software/lift.py").read()
    tree = ast.parse(src)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "lifter"][0]
    ns = {"np": np, "print": lambda *a, **k: None}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "software/lift.py", "exec"), ns)
    return ns["lifter"]


def main():
    import sklearn.preprocessing
    from oracle import ref_c as RC
    RC.build()
    wav = np.load(os.path.join(HERE, "f2bjrop1.npz"))["samples"]
    cep = RC.run(RC.RefConfig(nfilters=32, nceptrums=32), wav)[:400]
    lift = reference_lifter()(cep)
    sc = sklearn.preprocessing.scale(cep.T.astype(np.float64), axis=1)       # [32][frames] as in the script
    np.savez_compressed(os.path.join(HERE, "post_golden.npz"), cepstra=cep,
                        lift_f64=lift, lift_i16=lift.astype(np.int16),
                        scale_f64=sc.T.copy(), scale_i16=sc.astype(np.int16).T.copy())
    print("lift", lift.astype(np.int16)[0, :6], "scale", sc.astype(np.int16)[:4, 0])


if __name__ == "__main__":
    main()
