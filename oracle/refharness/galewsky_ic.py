"""Restated initial condition for config/galewsky.ini (case 8) -- TEST INFRASTRUCTURE ONLY.

At the pinned reference commit ``case_galewsky`` still assumes the old 2-D ``(ni, nj)`` layout
(``ni, nj = geom.lon.shape`` fails on the element-major ``(ny, nx, n^2)`` arrays:
wx_factory/init/shallow_water_test.py:289), so the reference cannot initialise this config.
Geometry, metric and RHS are fine for case 8; only the initial condition is restated here,
point by point with the SAME formulas (wx_factory/init/shallow_water_test.py:276-345 and
``eval_u_prime`` :12-26), over the flattened coordinate arrays, followed by the same
``geom.wind2contra(u, v)`` the other cases use (:271).

``install()`` monkey-patches the function into the (unmodified) reference modules in memory.
"""

import math

import numpy

gravity = 9.80616


def _u_prime(lat):
    u_max = 80.0
    phi0 = math.pi / 7.0
    phi1 = math.pi / 2.0 - phi0
    if lat < phi0 or lat > phi1:
        return 0.0
    e_n = math.exp(-4.0 / ((phi1 - phi0) ** 2))
    return u_max / e_n * math.exp(1.0 / ((lat - phi0) * (lat - phi1)))


def _height(lat, lon, earth_radius, rotation_speed):
    h0 = 10158.18617045463179
    h_hat = 120.0
    phi2 = math.pi / 4.0
    alpha = 1.0 / 3.0
    beta = 1.0 / 15.0
    n_int = max(int((lat + 0.5 * math.pi) / 1.0e-2), 1)
    lat_x = [-0.5 * math.pi + ((lat + 0.5 * math.pi) / n_int) * k for k in range(n_int + 1)]
    acc = 0.0
    for k in range(n_int):
        for m in (-1, 1):
            x = 0.5 * (lat_x[k + 1] + lat_x[k]) + m * math.sqrt(1.0 / 3.0) * 0.5 * (lat_x[k + 1] - lat_x[k])
            du = _u_prime(x)
            acc += (2.0 * earth_radius * rotation_speed * math.sin(x) + du * math.tan(x)) * du
    acc *= 0.5 * (lat_x[1] - lat_x[0])
    h = h0 - acc / gravity
    h += h_hat * math.cos(lat) * math.exp(-((lon / alpha) ** 2)) * math.exp(-(((phi2 - lat) / beta) ** 2))
    return h


def case_galewsky(geom, metric, param):
    lat = numpy.asarray(geom.lat)
    lon = numpy.asarray(geom.lon)
    shape = lat.shape
    h = numpy.empty(lat.size)
    u = numpy.empty(lat.size)
    for idx, (la, lo) in enumerate(zip(lat.ravel(), lon.ravel())):
        h[idx] = _height(float(la), float(lo), geom.earth_radius, geom.rotation_speed)
        u[idx] = _u_prime(float(la))  # v == 0, so both branches of :335-341 reduce to u_p
    u = u.reshape(shape)
    v = numpy.zeros(shape)
    u1, u2 = geom.wind2contra(u, v)
    return u1, u2, h.reshape(shape)


def install():
    import init.initialize
    import init.shallow_water_test

    init.shallow_water_test.case_galewsky = case_galewsky
    init.initialize.case_galewsky = case_galewsky
