"""N4: the dust-emission source term of the dust-heating iterations, effective-grain mode (set_dust_emission / calc_dens_dustem /
calc_t_dust_equil / zbrent_tdust, sed_routines.f90:249-276, 377-430, 846-874, 993-1078).  Oracle pins (CPU) and GPU parity."""
import numpy as np
import pytest

H, C0, KB, PC = np.float64(np.float32(6.62606957E-34)), np.float64(np.float32(2.99792458E8)), np.float64(np.float32(1.3806488E-23)), np.float64(np.float32(3.08567758E16))
PI = 2 * np.float64(np.float32(1.5707963705062866))


def planck(t, lam):
    return 2 * H * C0 ** 2 / lam ** 5 / (np.exp(H * C0 / (lam * KB * t)) - 1)


def inputs(scene, seed=5, ls=9, ld=40):
    from dartray_b200 import scenes
    rng = np.random.default_rng(seed)
    N = scene.grid.ncell
    lam_s = np.geomspace(0.1e-6, 3e-6, ls)
    lam_d = np.geomspace(3e-6, 1000e-6, ld)
    lam = np.concatenate([lam_s, lam_d])
    dl_s = scenes.delta_lambda_bins(lam)[:ls]          # delta_lambda_bin_stars: bins of the stellar part of the full list
    dl_d = scenes.delta_lambda_bins(lam)[ls:]
    kabs = 1e-25 * (lam / 1e-6) ** -1.5                 # m^2 per H, a smooth emissivity law
    dens_ref = np.where(scene.grid.cchild == -1, rng.uniform(0, 1, N), 0.0)
    dens_ref[rng.uniform(size=N) < 0.2] = 0.0           # dust-free leaves stay dark
    rf_s = rng.uniform(0.1, 10.0, (N, ls)) * 1e-2 * (lam_s / 1e-6) ** -2      # W/m/m^2
    rf_d = rng.uniform(0.1, 10.0, (N, ld)) * 1e-6
    return dict(kabs_stars=kabs[:ls], kabs_dust=kabs[ls:], lambda_si_dust=lam_d, dl_stars=dl_s, dl_dust=dl_d, kext_ref=3.2e-25,
                dens_ref=dens_ref, rf_stars=rf_s), rf_d


def test_oracle_equilibrium_temperature_balances(tiny_trusti):
    """The temperature Brent's method returns balances absorbed and emitted energy (independent numpy evaluation), the emissivity is
    4 pi kabs / kext_ref dens_ref B(T) parsec^2, dust-free and non-leaf cells stay dark."""
    from oracle import oracle as O
    orc = O.Oracle(tiny_trusti)
    a, rf_d = inputs(tiny_trusti)
    for itd, rfd in ((0, None), (2, rf_d)):
        dens, t = orc.dust_emission(**a, rf_dust=rfd, iterations_dustem=itd)
        on = (tiny_trusti.grid.cchild == -1) & (a["dens_ref"] != 0)
        assert np.all(t[~on] == 0) and np.all(dens[~on] == 0) and np.all(t[on] > 2.0) and np.all(t[on] < 2000.0)
        abs_en = (a["kabs_stars"] * a["rf_stars"] * a["dl_stars"]).sum(axis=1)
        if itd > 0:
            abs_en = abs_en + (a["kabs_dust"] * rfd * a["dl_dust"]).sum(axis=1)
        em = 4 * PI * (a["kabs_dust"] * planck(t[on, None], a["lambda_si_dust"]) * a["dl_dust"]).sum(axis=1)
        assert np.abs(em / abs_en[on] - 1).max() < 1e-10
        want = 4 * PI * a["kabs_dust"] / a["kext_ref"] * a["dens_ref"][on, None] * planck(t[on, None], a["lambda_si_dust"]) * PC ** 2
        assert np.abs(dens[on] / want - 1).max() < 1e-12


def test_oracle_grey_dust_follows_stefan_boltzmann(tiny_trusti):
    """Grey grains on a fine, wide wavelength grid: sum B dlambda -> sigma T^4 / pi, so T = (abs_en / (4 sigma kabs))^(1/4)."""
    from dartray_b200 import scenes
    from oracle import oracle as O
    orc = O.Oracle(tiny_trusti)
    N = tiny_trusti.grid.ncell
    lam = np.geomspace(0.05e-6, 5e-2, 1400)
    ls = 3
    dl = scenes.delta_lambda_bins(lam)
    sigma = 5.670374e-8
    t_want = 60.0
    abs_en = 4 * sigma * t_want ** 4                      # kabs = 1
    rf_s = np.zeros((N, ls)); rf_s[:, 0] = abs_en / dl[0]
    dens_ref = np.where(tiny_trusti.grid.cchild == -1, 1.0, 0.0)
    dens, t = orc.dust_emission(np.ones(ls), np.ones(len(lam) - ls), lam[ls:], dl[:ls], dl[ls:], 1.0, dens_ref, rf_s)
    on = dens_ref > 0
    assert np.abs(t[on] / t_want - 1).max() < 2e-3


@pytest.mark.gpu
def test_gpu_dust_emission_matches_oracle(tiny_galaxy):
    """Engine vs oracle: host radiation fields (first and later dust-heating iterations), then the resident form -- rf_dust read from
    the device's u_final, the result left on the device as lumcell = dens_stars * csize^3."""
    import copy
    from dartray_b200 import abi, engine, scenes
    from oracle import oracle as O
    from test_gpu_parity import assert_fields_close
    sc = tiny_galaxy
    L = sc.lnum
    orc = O.Oracle(sc)
    eng = engine.Engine(sc)
    a, rf_d = inputs(sc, ld=L)                      # the engine is configured for lnum wavelengths: they play the dust wavelengths
    for itd, rfd in ((0, None), (1, rf_d)):
        d_o, t_o = orc.dust_emission(**a, rf_dust=rfd, iterations_dustem=itd)
        d_g, t_g = eng.dust_emission(**a, rf_dust=rfd, iterations_dustem=itd)
        assert np.array_equal(t_g == 0, t_o == 0)
        assert_fields_close(t_g, t_o, f"t_dust, dust-heating iteration {itd}", rtol=1e-9)
        assert_fields_close(d_g, d_o, f"dens_stars, dust-heating iteration {itd}")
    # resident: u_final of the "dust" pass stays on the device
    u = np.random.default_rng(3).uniform(0.5, 2.0, (sc.grid.ncell, L)) * 1e30
    eng.upload(abi.ARR_U_FINAL, u)
    d_o, t_o = orc.dust_emission(**a, rf_dust=u / PC ** 3 * C0, iterations_dustem=3)
    d_g, t_g = eng.dust_emission(**a, rf_dust=None, iterations_dustem=3, to_lumcell=True)
    assert_fields_close(t_g, t_o, "t_dust (resident u_final)", rtol=1e-9)
    assert_fields_close(d_g, d_o, "dens_stars (resident u_final)")
    lum = np.zeros((sc.grid.ncell, L))
    eng.download(abi.ARR_LUMCELL, lum)
    assert_fields_close(lum, d_o * sc.grid.csize[:, None] ** 3, "lumcell = dens_stars * csize^3 left on the device")
    eng.close()
