set -x
python -m pytest tests -m gpu -x -q -rs > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2m_pytest.log; tail -6 gpurun_out/r2m_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')"
