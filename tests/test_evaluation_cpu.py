"""CPU: the host half of the evaluation tail (variable bookkeeping and the netCDF writer of dlwp_benchmark_b200.evaluation)."""
import numpy as np
import torch

from dlwp_benchmark_b200 import evaluation as ev

# the prognostic variables of configs/data/weatherbench.yaml:34-42 (t2m, u10, v10 without levels; t at 850; z at four levels)
VARS = {"t2m": [], "u10": [], "v10": [], "t": [850], "z": [300, 500, 700, 1000]}


def test_variable_order_and_statistics_follow_the_reference_loop():
    assert ev.variable_names(VARS) == ["t2m", "u10", "v10", "t850", "z300", "z500", "z700", "z1000"]
    stats = {"t2m": {"mean": 278.0, "std": 21.0}, "u10": {"mean": 0.0, "std": 5.5}, "v10": {"mean": 0.2, "std": 4.7},
             "t": {"level": {850: {"mean": 274.0, "std": 15.0}}},
             "z": {"level": {l: {"mean": 10.0 * l, "std": 100.0 + l} for l in (300, 500, 700, 1000)}}}
    ms = ev.variable_stats(VARS, stats)
    assert ms[0] == (278.0, 21.0) and ms[3] == (274.0, 15.0) and ms[7] == (10000.0, 1100.0) and len(ms) == 8


def test_forecast_file_round_trips(tmp_path):
    from scipy.io import netcdf_file
    torch.manual_seed(0)
    a = torch.randn(3, 2, 4, 8, 16)      # [C, B, T, H, W]
    path = ev.write_forecast(str(tmp_path / "outputs.nc"), a, ["t2m", "z500", "u10"], init_hours=[0.0, 24.0], step_hours=6.0,
                             attrs=[{"mean": 278.0, "std": 21.0}, {"mean": 5.4e4, "std": 3.3e3}, {"mean": 0.0, "std": 5.5}])
    with netcdf_file(path, "r", mmap=False) as f:
        assert f.variables["z500"].dimensions == ("sample", "time", "lat", "lon")
        assert np.array_equal(f.variables["z500"][:], a[1].numpy())
        assert np.allclose(f.variables["time"][:], [6.0, 12.0, 18.0, 24.0])
        lat = f.variables["lat"][:]
        assert np.isclose(lat[0], -90 + 180 / 8 / 2) and np.isclose(lat[-1], 90 - 180 / 8 / 2)      # cell centres
        assert abs(f.variables["t2m"].std - 21.0) < 1e-9
    # the initial-conditions file has no time axis; the fork's hard-coded 1-degree coordinates on request
    path = ev.write_forecast(str(tmp_path / "inits.nc"), a[:, :, 0], ["t2m", "z500", "u10"], fork_coordinates=True)
    with netcdf_file(path, "r", mmap=False) as f:
        assert f.variables["u10"].dimensions == ("sample", "lat", "lon")
        assert np.allclose(f.variables["lat"][:3], [-90.0, -89.0, -88.0])
