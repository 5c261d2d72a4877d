"""Host-side pieces of the command-line driver: the
MatrixMarket reader/writer against scipy.io (an independent implementation of
the same format) and the parameter-file reader against the Teuchos XML subset
the reference's main.cpp loads (/root/reference/src/main.cpp:55-61,117,143-151).
No GPU involved."""
import os

import numpy as np
import pytest
import scipy.io
import scipy.sparse as sp

from rails_b200 import api, problems

pytestmark = pytest.mark.filterwarnings("ignore::DeprecationWarning")


def _csr(rp, ci, v, shape):
    return sp.csr_matrix((v, ci, rp), shape=shape)


def test_read_crs_matches_scipy(tmp_path):
    rs = np.random.RandomState(3)
    A = sp.random(57, 57, density=0.08, random_state=rs, format="coo")
    A.data = rs.uniform(-1, 1, A.data.size)
    path = str(tmp_path / "A.mtx")
    scipy.io.mmwrite(path, A, precision=17)
    rp, ci, v, gr, nc = api.mm_read_crs(path)
    assert (gr, nc) == (57, 57)
    ref = A.tocsr()
    ref.sort_indices()
    assert np.array_equal(rp, ref.indptr)
    assert np.array_equal(ci, ref.indices)
    assert np.array_equal(v, ref.data)            # 17 digits: bit-exact


def test_read_crs_row_block_and_duplicates(tmp_path):
    path = str(tmp_path / "D.mtx")
    with open(path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real general\n% a comment\n\n"
                "4 3 6\n1 1 1.5\n3 2 2.0\n3 2 0.25\n4 3 -1\n2 1 7\n3 1 4\n")
    rp, ci, v, gr, nc = api.mm_read_crs(path)
    assert (gr, nc) == (4, 3)
    dense = _csr(rp, ci, v, (4, 3)).toarray()
    want = np.array([[1.5, 0, 0], [7, 0, 0], [4, 2.25, 0], [0, 0, -1.0]])
    assert np.array_equal(dense, want)            # the entry given twice is summed
    # rows [2, 4) only, global column indices kept
    rp2, ci2, v2, gr2, _ = api.mm_read_crs(path, 2, 4)
    assert gr2 == 4 and rp2.size == 3
    assert np.array_equal(_csr(rp2, ci2, v2, (2, 3)).toarray(), want[2:4])


def test_read_symmetric_pattern_integer(tmp_path):
    path = str(tmp_path / "S.mtx")
    with open(path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate integer symmetric\n3 3 4\n1 1 2\n2 1 -1\n3 2 5\n3 3 1\n")
    rp, ci, v, _, _ = api.mm_read_crs(path)
    got = _csr(rp, ci, v, (3, 3)).toarray()
    assert np.array_equal(got, scipy.io.mmread(path).toarray())
    path = str(tmp_path / "K.mtx")
    with open(path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate real skew-symmetric\n3 3 2\n2 1 1.5\n3 1 -2\n")
    rp, ci, v, _, _ = api.mm_read_crs(path)
    assert np.array_equal(_csr(rp, ci, v, (3, 3)).toarray(), scipy.io.mmread(path).toarray())
    path = str(tmp_path / "P.mtx")
    with open(path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate pattern general\n2 3 3\n1 1\n2 3\n1 2\n")
    rp, ci, v, _, nc = api.mm_read_crs(path)
    assert np.array_equal(_csr(rp, ci, v, (2, nc)).toarray(), np.array([[1.0, 1, 0], [0, 0, 1]]))


def test_dense_round_trip_is_bit_exact(tmp_path):
    rs = np.random.RandomState(5)
    V = rs.standard_normal((41, 6)) * 10.0 ** rs.randint(-12, 12, (41, 6))
    path = str(tmp_path / "V.mtx")
    api.mm_write_dense(path, V, "RAILS solution factor V")
    head = open(path).readline().strip()
    # the banner EpetraExt::MultiVectorToMatrixMarketFile writes (src/main.cpp:126-129)
    assert head == "%%MatrixMarket matrix array real general"
    assert np.array_equal(api.mm_read_dense(path), V)
    assert np.array_equal(scipy.io.mmread(path), V)       # scipy agrees on the layout
    # a file scipy wrote reads back the same through our reader
    path2 = str(tmp_path / "W.mtx")
    scipy.io.mmwrite(path2, V, precision=17)
    assert np.array_equal(api.mm_read_dense(path2), V)
    # empty block (a converged warm start can return zero eigenvectors)
    api.mm_write_dense(str(tmp_path / "E.mtx"), np.zeros((5, 0)))
    assert api.mm_read_dense(str(tmp_path / "E.mtx")).shape == (5, 0)


def test_crs_write_round_trip_and_generator(tmp_path):
    rp, ci, v = problems.laplacian(6, 2)
    path = str(tmp_path / "L.mtx")
    api.mm_write_crs(path, rp, ci, v, 36)
    A = scipy.io.mmread(path).tocsr()
    assert np.array_equal(A.toarray(), _csr(rp, ci, v, (36, 36)).toarray())
    rp2, ci2, v2, _, _ = api.mm_read_crs(path)
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2) and np.array_equal(v, v2)
    # a row block written with its global row numbers
    rb, cb, vb = problems.laplacian(6, 2, 12, 30)
    api.mm_write_crs(path, rb, cb, vb, 36, global_rows=36, row_begin=12)
    rp3, ci3, v3, _, _ = api.mm_read_crs(path, 12, 30)
    assert np.array_equal(rb, rp3) and np.array_equal(cb, ci3) and np.array_equal(vb, v3)


def test_reader_errors_are_reported(tmp_path):
    with pytest.raises(api.RailsError, match="cannot open"):
        api.mm_read_crs(str(tmp_path / "missing.mtx"))
    bad = str(tmp_path / "bad.mtx")
    open(bad, "w").write("%%MatrixMarket matrix coordinate real general\n2 2 2\n1 1 1.0\n")
    with pytest.raises(api.RailsError, match="malformed or missing"):
        api.mm_read_crs(bad)
    open(bad, "w").write("%%MatrixMarket matrix coordinate real general\n2 2 1\n3 1 1.0\n")
    with pytest.raises(api.RailsError, match="outside"):
        api.mm_read_crs(bad)
    open(bad, "w").write("%%MatrixMarket matrix array real general\n2 2\n1\n2\n3\n4\n")
    with pytest.raises(api.RailsError, match="coordinate"):
        api.mm_read_crs(bad)
    open(bad, "w").write("hello\n")
    with pytest.raises(api.RailsError, match="not a MatrixMarket"):
        api.mm_read_dense(bad)
    open(bad, "w").write("%%MatrixMarket matrix coordinate complex general\n1 1 1\n1 1 1 0\n")
    with pytest.raises(api.RailsError, match="complex"):
        api.mm_read_crs(bad)


XML = """<?xml version="1.0"?>
<!-- the layout the reference's main.cpp reads -->
<ParameterList name="RAILS">
  <ParameterList name="Lyapunov Solver">
    <Parameter name="Maximum iterations" type="int" value="500"/>
    <Parameter name="Tolerance" type="double" value="1e-4"/>
    <Parameter name="Minimize solution space" type="bool" value="false"/>
    <Parameter name="Expand size" type="int" value="3" />
    <Parameter name="Note" type="string" value="a &lt;b&gt; &amp; c"/>
  </ParameterList>
  <ParameterList name="Eigenvalue Solver">
    <Parameter name="Number of Eigenvalues" type="int" value="7"/>
    <Parameter name="Convergence Tolerance" type="double" value="1e-8"/>
    <ParameterList name="Nested"/>
  </ParameterList>
  <Parameter name="Top" type="double" value="2.5"/>
</ParameterList>
"""

INI = """# same content as an INI file
Top = 2.5

[Lyapunov Solver]
Maximum iterations = 500
Tolerance = 1e-4
Minimize solution space = false
Expand size = 3
; comment
[Eigenvalue Solver]
Number of Eigenvalues = 7
Convergence Tolerance = "1e-8"
"""


@pytest.mark.parametrize("text,name", [(XML, "p.xml"), (INI, "p.ini")])
def test_parameter_files(tmp_path, text, name):
    path = str(tmp_path / name)
    open(path, "w").write(text)
    assert api.params_get(path, "Lyapunov Solver", "Maximum iterations") == 500
    assert api.params_get(path, "Lyapunov Solver", "Tolerance") == 1e-4
    assert api.params_get(path, "Lyapunov Solver", "Minimize solution space") == 0.0
    assert api.params_get(path, "Lyapunov Solver", "Expand size") == 3
    assert api.params_get(path, "Eigenvalue Solver", "Number of Eigenvalues") == 7
    assert api.params_get(path, "Eigenvalue Solver", "Convergence Tolerance") == 1e-8
    assert api.params_get(path, "", "Top") == 2.5
    assert api.params_get(path, "Lyapunov Solver", "Restart size", -1) == -1
    assert api.params_get(path, "No such list", "Tolerance", "absent") == "absent"
    # re-emitted as XML and read again: same values
    again = str(tmp_path / "again.xml")
    open(again, "w").write(api.params_to_xml(path))
    assert api.params_get(again, "Lyapunov Solver", "Tolerance") == 1e-4
    assert api.params_get(again, "Eigenvalue Solver", "Number of Eigenvalues") == 7


def test_parameter_file_errors(tmp_path):
    path = str(tmp_path / "broken.xml")
    open(path, "w").write("<ParameterList name='a'><ParameterList name='b'></ParameterList>")
    with pytest.raises(api.RailsError, match="not closed"):
        api.params_get(path, "", "x")
    open(path, "w").write("<ParameterList><Parameter name='x'/></ParameterList>")
    with pytest.raises(api.RailsError, match="name and value"):
        api.params_get(path, "", "x")
    open(path, "w").write("<Other/>")
    with pytest.raises(api.RailsError, match="unexpected element"):
        api.params_get(path, "", "x")
    with pytest.raises(api.RailsError, match="cannot open"):
        api.params_get(str(tmp_path / "nope.xml"), "", "x")


def test_cli_binary_exists_and_refuses_without_gpu(tmp_path):
    import subprocess
    exe = os.path.join(os.path.dirname(api.HERE), "rails_b200", "rails_cuda_main")
    assert os.path.exists(exe), "run `python -m rails_b200.build`"
    r = subprocess.run([exe, "--help"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0 and "usage" in r.stderr
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if not has_gpu:
        r = subprocess.run([exe, "--dir", str(tmp_path)], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                           text=True)
        assert r.returncode == 3 and "no CPU fallback" in r.stderr
