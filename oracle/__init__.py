"""CPU oracle for the CurveCloudNet per-scan hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``curvecloudnet_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU
baseline legs use it, and only as the checker / the timed CPU baseline.

Layout
------
tp_ops.py       restatement of the third-party arithmetic the reference calls but
                does not vendor (pytorch3d @stable, torch_scatter 2.1.1,
                torch_geometric 2.3.0) plus the vendored FRNN result contract.
tp_nn.py        torch_geometric.nn.MLP / Batch / MessagePassing look-alikes with the
                same parameter names.
ccn_point_ops.py, ccn_modules.py, ccn_model.py
                restatement of /root/reference/src/models/** (file:line cited per
                function).
shims/          import shims that expose tp_ops under the third-party package
                names so that the REAL reference python sources can be imported in
                the build container (make_golden.py) to pin the restatement.
_ref/           build output of the one reference C++ file that compiles stand-alone
                (FRNN bruteforce_cpu.cpp); see Makefile.

Parity status: the reference's own python (base.py, modules/*, point_ops.py,
fps_ops.py) is PINNED by golden vectors generated from the real sources run over
the shims (tests/golden/*, oracle/make_golden.py).  The third-party arithmetic
underneath the shims (pytorch3d FPS/ball_query/knn_points, torch_scatter, PyG MLP)
is a restatement from their published behaviour and is "parity unpinned" except
the fixed-radius kNN, which is cross-checked against the compiled reference
FRNNBruteForceCPU.
"""
