"""CPU: the .param loader (host logic) on CONCEPT-style parameter text written like the
reference's files (forward references, units, non-ASCII names, string grid expressions)."""
import numpy as np
import os
import pytest

from nbody_b200 import params

TEXT = '''
_size = 32
_tag = f"NNPM32_As{primordial_spectrum['A_s']:.2e}"
initial_conditions = {'species': 'matter', 'N': _size**3}
output_dirs = {'snapshot': f'{path.output_dir}/{_tag}/snapshots'}
boxsize = 500*Mpc
potential_options = {
    'gridsize': {'global': {_size}, 'particles': {'gravity': {'p3m': ['cbrt(N)', 'cbrt(N)']}}},
    'interlace': {'gravity': {'p3m': (True, True)}},
    'deconvolve': {'gravity': {'p3m': (True, False)}},
}
select_forces = {'matter': {'gravity': 'p3m'}}
H0 = 67.36*km/(s*Mpc)
Ωb = 0.02237/h**2
Ωcdm = 0.12/h**2
a_begin = 0.02
autosave_interval = 6*hr
primordial_spectrum = {'A_s': 2.105e-9, 'n_s': 0.9649}
broken = undefined_name + 1
'''


def test_forward_references_units_and_unresolved():
    v, unresolved = params.load_param(TEXT, is_text=True)
    assert v["_tag"] == "NNPM32_As2.10e-09" or v["_tag"] == "NNPM32_As2.11e-09"
    assert abs(v["H0"] - 0.0688899) < 1e-6                       # Gyr^-1, as in the snapshot attrs
    assert abs(v["h"] - 0.6736) < 1e-12
    assert abs(v["Ωb"] - 0.02237 / 0.6736 ** 2) < 1e-12
    assert len(unresolved) == 1 and "undefined_name" in unresolved[0]


def test_force_config_fields_and_mesh_override():
    c = params.force_config(TEXT, physical_units=True)
    assert c.method == "p3m" and c.npart == 32768 and c.gridsize == 32 and c.boxsize == 500.0
    assert c.interlace == (True, True) and c.deconvolve == (True, False) and c.deconv_order == 1
    assert abs(c.mass - 15072.97) < 0.01                         # snapshot attr `mass`
    assert abs(c.G - 4.4985e-5) < 1e-8
    assert abs(c.r_s - 1.25 * 500 / 32) < 1e-12 and abs(c.r_cut - 4.5 * c.r_s) < 1e-12
    assert params.force_config(TEXT, mesh=64).gridsize == 64


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="reference tree not present")
def test_reference_param_files_load():
    for f, meth, n, L in (("Different_Sizes/Giant.param", "pm", 256 ** 3, 1000.0),
                          ("TheFeedingPipeline/NN.param", "p3m", 32 ** 3, 500.0),
                          ("SnapVsPK/Uncertain.param", "pm", 64 ** 3, 1000.0)):
        c = params.force_config(os.path.join("/root/reference", f))
        assert (c.method, c.npart, c.boxsize) == (meth, n, L)
        assert c.gridsize == round(n ** (1 / 3)) and c.interlace == (True, True)
