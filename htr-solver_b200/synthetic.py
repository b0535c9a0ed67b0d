"""Deterministic synthetic inputs for the RHS hot path (numpy, host side).

Taylor-Green 3-D vortex primitives as the reference initialises them
(src/prometeo_init.rg:134-158), on a tri-periodic box with cell centres (i+1/2) L/N, uniform or
smoothly stretched metrics, all-standard node types (periodic directions,
src/prometeo_metric.rg:142-148).  Only primitives are produced here; density, transport
properties and the conserved vector come from the UpdatePropertiesFromPrimitive /
UpdateConservedFromPrimitive kernels (or, in the tests, from the oracle).
"""
import numpy as np

# reference input files these parameter sets come from
CONSTPROP_3DPERIODIC = {"type": "ConstPropMix", "gasConstant": 1.0, "gamma": 1.4, "prandtl": 0.71,
                        "viscosityModel": {"type": "Constant", "Visc": 0.0}}          # solverTests/3DPeriodic_TENO/base.json
CONSTPROP_VISCOUS = {"type": "ConstPropMix", "gasConstant": 1.0, "gamma": 1.4, "prandtl": 0.71,
                     "viscosityModel": {"type": "Constant", "Visc": 5.0e-3}}
AIRMIX_3DPERIODIC = {"type": "AirMix", "LRef": 1.0, "TRef": 500.0, "PRef": 100000.0,
                     "XiRef": {"Species": [{"Name": "N2", "MolarFrac": 0.79}, {"Name": "O2", "MolarFrac": 0.21}]}}  # solverTests/3DPeriodic_Air/base.json
AIR_SPECIES = ["N2", "O2", "NO", "N", "O"]
# one-step methane oxidation (src/Mixtures/CH41StMix.hpp); reference state as in unitTests/mixTest/CH41StMix.json
CH41ST_3DPERIODIC = {"type": "CH41StMix", "LRef": 1.0, "TRef": 300.0, "PRef": 101325.0,
                     "XiRef": {"Species": [{"Name": "O2", "MolarFrac": 1.0}]}}
SPECIES = {"AirMix": AIR_SPECIES, "CH41StMix": ["CH4", "O2", "CO2", "H2O"]}


def grid_nodes(n, origin, width, kind="Uniform", stretching=1.0):
    """Node positions of one direction as the reference places them (src/prometeo_grid.rg:123-230: InitializeNodeGrid +
    transform_uniform_to_nonuniform): n+1 nodes on [origin, origin+width], uniform or mapped by the Tanh family."""
    xmin, xmax = origin, origin + width
    u = xmin + np.arange(n + 1) * ((xmax - xmin) / n)
    lin = lambda x, x0, x1, y0, y1: y0 + (x - x0) * (y1 - y0) / (x1 - x0)
    if kind == "Uniform":
        return u
    if kind == "Tanh":
        s = lin(u, xmin, xmax, -1.0, 1.0)
        return lin(np.tanh(stretching * s) / np.tanh(stretching), -1.0, 1.0, xmin, xmax)
    if kind == "TanhMinus":
        s = lin(u, xmin, xmax, -1.0, 0.0)
        return lin(np.tanh(stretching * s) / np.tanh(stretching), -1.0, 0.0, xmin, xmax)
    if kind == "TanhPlus":
        s = lin(u, xmin, xmax, 0.0, 1.0)
        return lin(np.tanh(stretching * s) / np.tanh(stretching), 0.0, 1.0, xmin, xmax)
    raise ValueError(kind)


def cell_centers(nodes):
    """Interior cell centres (src/prometeo_grid.rg:313-316): the mean of the two nodes."""
    return 0.5 * (nodes[1:] + nodes[:-1])


def taylor_green(n, mix_cfg, p0=None, T0=None, U=None, width=2 * np.pi, stretch=0.0, species_wave=0.0,
                 front=0.0, dtype=np.float64, zslab=None, offset=(0, 0, 0), global_n=None, ygrid=None):
    """Host fields of an (nx,ny,nz)=n periodic box.  Returns dict name -> array [z][y][x](comp).

    stretch > 0      multiplies the metrics by 1 + stretch*sin(coordinate): a smooth non-uniform grid
    species_wave > 0 (multi-component only) adds smooth O2/NO/N/O mole-fraction waves
    front > 0        superposes a tanh temperature/pressure front of that relative strength in x
                     (steep gradients: exercises the nonlinear TENO weights and the shock sensor)
    """
    nx, ny, nz = (n, n, n) if np.isscalar(n) else n
    gx, gy, gz = (nx, ny, nz) if global_n is None else ((global_n,) * 3 if np.isscalar(global_n) else global_n)
    multi = mix_cfg["type"] != "ConstPropMix"
    names = SPECIES.get(mix_cfg["type"], ["MIX"])
    nspec = len(names)
    if p0 is None:
        p0, T0, U = (1.0, 1.0, 0.002634364032587426) if multi else (100.0, 100.0, 1.0)
    # (nx,ny,nz) is this block of a (gx,gy,gz) global box, starting at cell `offset`;
    # zslab=(k0,k1) restricts the output to local planes k0..k1-1 (slab-wise generation of big boxes)
    k0, k1 = (0, nz) if zslab is None else zslab
    x = (np.arange(nx) + offset[0] + 0.5) * width / gx
    y = (np.arange(ny) + offset[1] + 0.5) * width / gy
    z = (np.arange(k0, k1) + offset[2] + 0.5) * width / gz
    nz = k1 - k0
    if ygrid is not None:
        # y stretched like the reference's grids: cell centres and the three y metrics of the GLOBAL (periodic) direction come
        # from the caller (dict: centers, deta_e, deta_d, deta_s, each of length gy: the reference's InitializeMetricTask
        # evaluated through the oracle by tests/golden/gen_tanh_metrics.py, or a committed fixture of it)
        jy = (np.arange(ny) + offset[1]) % gy
        y = np.asarray(ygrid["centers"])[jy]
    Z, Y, X = np.meshgrid(z, y, x, indexing="ij")
    f = {}
    for d, (nn, c) in enumerate(((gx, X), (gy, Y), (gz, Z))):
        m = (nn / width) * (1.0 + stretch * np.sin(c + 0.3 * d))
        for kind in "eds":
            f[("dcsi_", "deta_", "dzet_")[d] + kind] = np.ascontiguousarray(m * (1.0 + 0.01 * stretch * "eds".index(kind)))
        f[("nType_x", "nType_y", "nType_z")[d]] = np.zeros((nz, ny, nx), dtype=np.int32)
    if ygrid is not None:
        for kind in "eds":
            f["deta_" + kind] = np.ascontiguousarray(np.broadcast_to(np.asarray(ygrid["deta_" + kind])[jy][None, :, None], (nz, ny, nx)))
    R = mix_cfg.get("gasConstant", 1.0)
    rho0 = p0 / (R * T0)
    vel = np.zeros((nz, ny, nx, 3))
    vel[..., 0] = U * np.sin(X) * np.cos(Y) * np.cos(Z)
    vel[..., 1] = -U * np.cos(X) * np.sin(Y) * np.cos(Z)
    P = p0 + rho0 * U ** 2 / 16.0 * (np.cos(2 * Z) + 2.0) * (np.cos(2 * X) + np.cos(2 * Y))
    T = P / (rho0 * R)
    if front > 0.0:
        s = np.tanh((X - 0.5 * width) * 8.0 / width * gx / 32.0) * np.cos(Y) * np.cos(Z)
        T = T * (1.0 + front * s)
        P = P * (1.0 + 0.5 * front * s)
    Xi = np.zeros((nz, ny, nx, nspec))
    if multi:
        ref = {s["Name"]: s["MolarFrac"] for s in mix_cfg["XiRef"]["Species"]}
        for i, s in enumerate(names):
            Xi[..., i] = ref.get(s, 1.0e-60)
        if species_wave > 0.0 and mix_cfg["type"] == "AirMix":
            w = species_wave
            Xi[..., 1] += w * 0.05 * np.sin(X) * np.cos(2 * Y)
            Xi[..., 2] = w * 0.01 * (1.1 + np.sin(2 * X + Z))
            Xi[..., 3] = w * 0.002 * (1.1 + np.cos(Y + 2 * Z))
            Xi[..., 4] = w * 0.005 * (1.1 + np.sin(X + Y + Z))
            Xi[..., 0] = 1.0 - Xi[..., 1:].sum(axis=-1)
        elif species_wave > 0.0:
            # any other mixture: smooth waves on every species but the dominant one of the reference state
            w = species_wave
            dom = int(np.argmax([ref.get(s, 0.0) for s in names]))
            k = 0
            for i in range(nspec):
                if i != dom:
                    Xi[..., i] = w * (0.04 / (1 + k)) * (1.1 + np.sin((1 + k % 2) * X + (k // 2) * Y + Z + 0.7 * k))
                    k += 1
            Xi[..., dom] = 1.0 - (Xi.sum(axis=-1) - Xi[..., dom])
    else:
        Xi[..., 0] = 1.0
    f["pressure"] = np.ascontiguousarray(P, dtype=dtype)
    f["temperature"] = np.ascontiguousarray(T, dtype=dtype)
    f["velocity"] = vel
    f["MolarFracs"] = Xi
    return f
