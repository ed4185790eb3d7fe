#!/usr/bin/env python3
"""Procedural .pbrt scenes for the configs of BASELINE.json.

    python scenes/gen_scenes.py <name> <out.pbrt> [key=value ...]

names:  cornell       C2: spectral Cornell box (diffuse + glass + metal), perspective camera
        cornell_lens  C3: the same scene through Camera "realistic" (double-Gauss lens)
        instanced     C4: N x N displaced grid mesh, instanced on a lattice
        zoo           one of every supported material / shape / light / camera feature (parity)
        envbox        open scene under Light "infinite" (env=const | env=map), SURVEY 8(f) rank 3
        manylights    an emissive 24 x 24 mesh = 1152 area lights under the spatial light distribution
        texbox        Texture "imagemap" Kd on matte / plastic (lens=R: thin-lens camera), 8(f) rank 3
        furnace_point / furnace_area   the known-answer scenes of the reference's
                      src/tests/analytic_scenes.cpp:68-165 written as scene files
keys:   res=<w>[x<h>] spp=<n> maxdepth=<n> strategy=<uniform|power|spatial> sampler=<halton|random>
        grid=<N> inst=<k> (instanced: k x k instances)
The files only use directives the reference's parser accepts, so the same file
feeds the reference CLI (oracle/_ref/pbrt) and the GPU host binary
(host/_build/pbrt_b200).
"""
import math
import os
import sys

import numpy as np

# 11-interface double-Gauss, 50 mm (curvature radius, thickness, eta, aperture
# diameter; mm).  The table of pbrt-v3-scenes' dgauss.50mm.dat as recalled in
# SURVEY.md 8d; oracle and GPU read the same file, so any consistent table works.
DGAUSS_50MM = """# D-GAUSS F/2 22deg HFOV (50 mm)
29.475 3.76 1.67 25.2
84.83 0.12 1 25.2
19.275 4.025 1.67 23
40.77 3.275 1.699 23
12.75 5.705 1 18
0 4.5 0 17.1
-14.495 1.18 1.603 17
40.77 6.065 1.658 20
-20.385 0.19 1 20
437.065 3.22 1.717 20
-39.73 5.0 1 20
"""

WHITE = "[400 .73 700 .73]"
RED = "[400 .05 580 .05 600 .65 700 .65]"
GREEN = "[400 .1 500 .1 520 .45 560 .45 580 .1 700 .1]"


def parse_res(v):
    if "x" in v:
        a, b = v.split("x")
        return int(a), int(b)
    return int(v), int(v)


def header(opts, camera, name):
    w, h = opts["res"]
    s = []
    s.append(camera)
    s.append(f'Film "image" "integer xresolution" [{w}] "integer yresolution" [{h}] '
             f'"string filename" "{name}.exr"')
    s.append(f'Sampler "{opts.get("sampler", "halton")}" "integer pixelsamples" [{opts["spp"]}]')
    integ = 'Integrator "path"'
    if "metadata" in opts:  # MetadataIntegrator: depth / material / mesh / coordinates
        integ = f'Integrator "metadata" "string strategy" "{opts["metadata"]}"'
    elif "bands" in opts:  # SpectralPathIntegrator: one path per wavelength band
        integ = f'Integrator "spectralpath" "integer numCABands" [{int(opts["bands"])}]'
    s.append(f'{integ} "integer maxdepth" [{opts["maxdepth"]}] '
             f'"string lightsamplestrategy" "{opts["strategy"]}"')
    return "\n".join(s) + "\n"


def quad(p0, p1, p2, p3, uv=None):
    pts = " ".join(f"{c:g}" for p in (p0, p1, p2, p3) for c in p)
    st = f' "float uv" [{" ".join(f"{c:g}" for c in uv)}]' if uv else ""
    return f'Shape "trianglemesh" "integer indices" [0 1 2 0 2 3] "point P" [{pts}]{st}\n'


def cornell_world():
    s = "WorldBegin\n"
    # emitter: two triangles facing -y, just below the ceiling -> 2 area lights
    s += 'AttributeBegin\n  Material "matte" "spectrum Kd" [400 0 700 0]\n'
    s += '  AreaLightSource "diffuse" "blackbody L" [6500 15]\n'
    s += "  " + quad((-.25, .99, -.25), (.25, .99, -.25), (.25, .99, .25), (-.25, .99, .25))
    s += "AttributeEnd\n"
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {WHITE}\n'
    s += "  " + quad((-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1))      # back
    s += "  " + quad((-1, -1, 1), (1, -1, 1), (1, -1, -1), (-1, -1, -1))      # floor
    s += "  " + quad((-1, 1, -1), (1, 1, -1), (1, 1, 1), (-1, 1, 1))          # ceiling
    s += "AttributeEnd\n"
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {RED}\n'
    s += "  " + quad((-1, -1, 1), (-1, -1, -1), (-1, 1, -1), (-1, 1, 1)) + "AttributeEnd\n"
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {GREEN}\n'
    s += "  " + quad((1, -1, -1), (1, -1, 1), (1, 1, 1), (1, 1, -1)) + "AttributeEnd\n"
    s += ('AttributeBegin\n  Material "glass" "float index" [1.5]\n  Translate .45 -.65 .3\n'
          '  Shape "sphere" "float radius" [.35]\nAttributeEnd\n')
    s += ('AttributeBegin\n  Material "metal" "float roughness" [.01]\n  Translate -.45 -.6 -.3\n'
          '  Shape "sphere" "float radius" [.4]\nAttributeEnd\n')
    s += "WorldEnd\n"
    return s


def uber_world():
    """The Cornell box with Material "uber" everywhere (src/materials/uber.cpp): diffuse + glossy
    walls, a partly transparent wall (opacity < 1: the pass-through SpecularTransmission lobe), a
    specular reflect + transmit sphere (Kr, Kt, index) and a default uber sphere."""
    s = "WorldBegin\n"
    s += 'AttributeBegin\n  Material "matte" "spectrum Kd" [400 0 700 0]\n'
    s += '  AreaLightSource "diffuse" "blackbody L" [6500 15]\n'
    s += "  " + quad((-.25, .99, -.25), (.25, .99, -.25), (.25, .99, .25), (-.25, .99, .25))
    s += "AttributeEnd\n"
    s += (f'AttributeBegin\n  Material "uber" "spectrum Kd" {WHITE} "spectrum Ks" [400 .2 700 .1] '
          '"float roughness" [.2]\n')
    s += "  " + quad((-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1))      # back
    s += "  " + quad((-1, -1, 1), (1, -1, 1), (1, -1, -1), (-1, -1, -1))      # floor
    s += "  " + quad((-1, 1, -1), (1, 1, -1), (1, 1, 1), (-1, 1, 1))          # ceiling
    s += "AttributeEnd\n"
    # a matte wall behind the partly transparent one, so that what passes through lands somewhere
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {GREEN}\n'
    s += "  " + quad((-1.4, -1, 1), (-1.4, -1, -1), (-1.4, 1, -1), (-1.4, 1, 1)) + "AttributeEnd\n"
    s += (f'AttributeBegin\n  Material "uber" "spectrum Kd" {RED} "spectrum Ks" [400 0 700 0] '
          '"spectrum opacity" [400 .7 550 .5 700 .3]\n')
    s += "  " + quad((-1, -1, 1), (-1, -1, -1), (-1, 1, -1), (-1, 1, 1)) + "AttributeEnd\n"
    s += (f'AttributeBegin\n  Material "uber" "spectrum Kd" {GREEN} "float uroughness" [.05] '
          '"float vroughness" [.3] "bool remaproughness" "false"\n')
    s += "  " + quad((1, -1, -1), (1, -1, 1), (1, 1, 1), (1, 1, -1)) + "AttributeEnd\n"
    s += ('AttributeBegin\n  Material "uber" "spectrum Kd" [400 0 700 0] "spectrum Ks" [400 0 700 0] '
          '"spectrum Kr" [400 .9 700 .7] "spectrum Kt" [400 .6 700 .9] "float index" [1.5]\n'
          '  Translate .45 -.65 .3\n  Shape "sphere" "float radius" [.35]\nAttributeEnd\n')
    # all five lobes at once: pass-through, diffuse, glossy, specular reflection and transmission
    s += ('AttributeBegin\n  Material "uber" "spectrum Kr" [400 .2 700 .3] "spectrum Kt" [400 .3 700 .2] '
          '"spectrum opacity" [400 .8 700 .9]\n  Translate -.45 -.6 -.3\n'
          '  Shape "sphere" "float radius" [.4]\nAttributeEnd\n')
    s += "WorldEnd\n"
    return s


def lobes_world():
    """Materials whose lobes go beyond the first five: "substrate" (FresnelBlend), "translucent"
    (Lambertian and microfacet transmission) and rough "glass" (microfacet reflection +
    transmission), in the Cornell box."""
    s = "WorldBegin\n"
    s += 'AttributeBegin\n  Material "matte" "spectrum Kd" [400 0 700 0]\n'
    s += '  AreaLightSource "diffuse" "blackbody L" [6500 15]\n'
    s += "  " + quad((-.25, .99, -.25), (.25, .99, -.25), (.25, .99, .25), (-.25, .99, .25))
    s += "AttributeEnd\n"
    s += (f'AttributeBegin\n  Material "substrate" "spectrum Kd" {WHITE} "spectrum Ks" [400 .1 700 .3] '
          '"float uroughness" [.1] "float vroughness" [.25]\n')
    s += "  " + quad((-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1))      # back
    s += "AttributeEnd\n"
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {WHITE}\n'
    s += "  " + quad((-1, -1, 1), (1, -1, 1), (1, -1, -1), (-1, -1, -1))      # floor
    s += "  " + quad((-1, 1, -1), (1, 1, -1), (1, 1, 1), (-1, 1, 1))          # ceiling
    s += "AttributeEnd\n"
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {RED}\n'
    s += "  " + quad((-1, -1, 1), (-1, -1, -1), (-1, 1, -1), (-1, 1, 1)) + "AttributeEnd\n"
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {GREEN}\n'
    s += "  " + quad((1, -1, -1), (1, -1, 1), (1, 1, 1), (1, 1, -1)) + "AttributeEnd\n"
    # a translucent panel standing in the box
    s += ('AttributeBegin\n  Material "translucent" "spectrum Kd" [400 .5 700 .6] "spectrum Ks" [400 .3 700 .2] '
          '"spectrum reflect" [400 .6 700 .5] "spectrum transmit" [400 .4 700 .7] "float roughness" [.2]\n')
    s += "  " + quad((-.9, -1, .2), (-.1, -1, -.5), (-.1, .3, -.5), (-.9, .3, .2)) + "AttributeEnd\n"
    s += ('AttributeBegin\n  Material "glass" "float index" [1.5] "float uroughness" [.1] "float vroughness" [.2]\n'
          '  Translate .45 -.65 .3\n  Shape "sphere" "float radius" [.35]\nAttributeEnd\n')
    s += ('AttributeBegin\n  Material "substrate" "spectrum Kd" [400 .6 700 .2] "spectrum Ks" [400 .5 700 .5] '
          '"float uroughness" [.05] "float vroughness" [.05] "bool remaproughness" "false"\n'
          '  Translate .2 -.7 -.45\n  Shape "sphere" "float radius" [.3]\nAttributeEnd\n')
    s += "WorldEnd\n"
    return s


def lobes_box(opts, out):
    cam = 'LookAt 0 0 3.8  0 0 0  0 1 0\nCamera "perspective" "float fov" [39]'
    return header(opts, cam, os.path.splitext(os.path.basename(out))[0]) + lobes_world()


def uber_box(opts, out):
    cam = 'LookAt 0 0 3.8  0 0 0  0 1 0\nCamera "perspective" "float fov" [39]'
    return header(opts, cam, os.path.splitext(os.path.basename(out))[0]) + uber_world()


def write_pfm(path, img):
    """img [H][W][3] float32, row 0 = top; PFM stores scanlines bottom to top."""
    h, w, _ = img.shape
    with open(path, "wb") as f:
        f.write(f"PF\n{w} {h}\n-1.0\n".encode())
        f.write(np.ascontiguousarray(img[::-1].astype("<f4")).tobytes())


def envbox(opts, out):
    """Open scene under an InfiniteAreaLight (src/lights/infinite.cpp): a floor, a matte, a glass and
    a metal sphere and a small area light, lit by a constant sky (env=const) or by a procedural
    12 x 6 lat-long map (env=map; not a power of two, so the MIPMap resamples it) with a sun."""
    name = os.path.splitext(os.path.basename(out))[0]
    cam = 'LookAt 0 1.2 4.2  0 -.2 0  0 1 0\nCamera "perspective" "float fov" [42]'
    s = header(opts, cam, name) + "WorldBegin\n"
    s += "AttributeBegin\n  Rotate -90 1 0 0\n  Rotate 25 0 0 1\n"
    if opts.get("env", "const") == "map":
        w, h = 12, 6
        v, u = np.meshgrid((np.arange(h) + .5) / h, (np.arange(w) + .5) / w, indexing="ij")
        img = np.zeros((h, w, 3), np.float32)
        img[..., 0] = .25 + .5 * u                     # horizon colours vary with azimuth
        img[..., 1] = .35 + .3 * np.sin(3.1 * u) ** 2
        img[..., 2] = .9 - .6 * v                      # bluer towards the zenith
        img[v > .55] *= .15                            # dark ground half
        img[1, 3] = (40., 36., 28.)                    # a sun
        write_pfm(os.path.join(os.path.dirname(os.path.abspath(out)), name + "_env.pfm"), img)
        s += f'  LightSource "infinite" "string mapname" "{name}_env.pfm" "rgb L" [1.5 1.5 1.5]\n'
    else:
        s += '  LightSource "infinite" "blackbody L" [6000 2.5]\n'
    s += "AttributeEnd\n"
    s += 'AttributeBegin\n  AreaLightSource "diffuse" "blackbody L" [3200 20]\n'
    s += "  " + quad((-.2, 1.4, -.2), (.2, 1.4, -.2), (.2, 1.4, .2), (-.2, 1.4, .2)) + "AttributeEnd\n"
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {WHITE}\n'
    s += "  " + quad((-3, -1, 3), (3, -1, 3), (3, -1, -3), (-3, -1, -3)) + "AttributeEnd\n"
    s += (f'AttributeBegin\n  Material "matte" "spectrum Kd" {RED}\n  Translate -1 -.55 0\n'
          '  Shape "sphere" "float radius" [.45]\nAttributeEnd\n')
    s += ('AttributeBegin\n  Material "glass" "float index" [1.5]\n  Translate 0 -.6 .5\n'
          '  Shape "sphere" "float radius" [.4]\nAttributeEnd\n')
    s += ('AttributeBegin\n  Material "metal" "float roughness" [.01]\n  Translate 1 -.5 -.2\n'
          '  Shape "sphere" "float radius" [.5]\nAttributeEnd\n')
    return s + "WorldEnd\n"


def texbox(opts, out):
    """Image textures (src/textures/imagemap.h) as the Kd of matte and plastic: a 20 x 12 procedural
    PFM (not a power of two, so the MIPMap resamples it) looked up with the EWA filter (floor, tiled
    by uscale / vscale, wrap "repeat"), trilinearly (a plastic sphere, wrap "clamp") and with wrap
    "black" (a matte sphere: outside [0,1]^2 the texture, hence the BSDF, is black).  lens=R gives
    the camera a thin lens (the other branch of PerspectiveCamera::GenerateRayDifferential),
    camera=realistic the double-Gauss RealisticCamera (Camera::GenerateRayDifferential)."""
    name = os.path.splitext(os.path.basename(out))[0]
    cam = 'LookAt 0 1.1 4.0  0 -.3 0  0 1 0\nCamera "perspective" "float fov" [40]'
    if float(opts.get("lens", 0)) > 0:
        cam += f' "float lensradius" [{float(opts["lens"]):g}] "float focaldistance" [4.2]'
    if opts.get("camera") == "realistic":  # Camera::GenerateRayDifferential: offset rays traced through the lens
        with open(os.path.join(os.path.dirname(os.path.abspath(out)), "dgauss.50mm.dat"), "w") as f:
            f.write(DGAUSS_50MM)
        cam = ('LookAt 0 1.1 4.0  0 -.3 0  0 1 0\nCamera "realistic" "string lensfile" "dgauss.50mm.dat" '
               '"float aperturediameter" [8] "float focusdistance" [4.2] "bool simpleweighting" "true"')
        if opts.get("ca") == "true":
            cam += ' "bool chromaticAberrationEnabled" "true"'
    s = header(opts, cam, name) + "WorldBegin\n"
    w, h = 20, 12
    v, u = np.meshgrid((np.arange(h) + .5) / h, (np.arange(w) + .5) / w, indexing="ij")
    img = np.zeros((h, w, 3), np.float32)
    check = ((np.floor(u * 5) + np.floor(v * 3)) % 2).astype(np.float32)
    img[..., 0] = .15 + .7 * check
    img[..., 1] = .2 + .6 * u
    img[..., 2] = .75 - .6 * v * check
    img[5:7, 8:12] = 0.                                 # a black patch: no diffuse lobe there
    write_pfm(os.path.join(os.path.dirname(os.path.abspath(out)), name + "_tex.pfm"), img)
    tex = f'"string filename" "{name}_tex.pfm"'
    s += f'Texture "floor" "spectrum" "imagemap" {tex} "float uscale" [3] "float vscale" [2.5] "float udelta" [.25]\n'
    s += f'Texture "ball" "spectrum" "imagemap" {tex} "bool trilinear" "true" "string wrap" "clamp" "float uscale" [2]\n'
    s += f'Texture "patch" "spectrum" "imagemap" {tex} "string wrap" "black" "float uscale" [1.6] "float vscale" [1.6] "float vdelta" [-.3] "float maxanisotropy" [4]\n'
    s += 'AttributeBegin\n  AreaLightSource "diffuse" "blackbody L" [5500 18]\n'
    s += "  " + quad((-.3, 1.6, -.3), (.3, 1.6, -.3), (.3, 1.6, .3), (-.3, 1.6, .3)) + "AttributeEnd\n"
    s += 'LightSource "point" "point from" [-2 1.5 2] "blackbody I" [4000 6]\n'
    floor_mat = 'Material "matte" "texture Kd" "floor"'
    if opts.get("floor") == "uber":  # UberMaterial with a textured Kd: Lambert + Trowbridge-Reitz lobes
        floor_mat = 'Material "uber" "texture Kd" "floor" "spectrum Ks" [400 .2 700 .2] "float roughness" [.15]'
    s += f'AttributeBegin\n  {floor_mat}\n'
    s += "  " + quad((-3, -1, 3), (3, -1, 3), (3, -1, -6), (-3, -1, -6), uv=(0, 0, 1, 0, 1, 1, 0, 1)) + "AttributeEnd\n"
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {WHITE}\n'
    s += "  " + quad((-3, -1, -6), (3, -1, -6), (3, 3, -6), (-3, 3, -6)) + "AttributeEnd\n"
    s += ('AttributeBegin\n  Material "plastic" "texture Kd" "ball" "spectrum Ks" [400 .3 700 .3] "float roughness" [.08]\n'
          '  Translate -.9 -.45 .2\n  Rotate 30 0 1 0\n  Shape "sphere" "float radius" [.55]\nAttributeEnd\n')
    s += ('AttributeBegin\n  Material "matte" "texture Kd" "patch"\n  Translate .9 -.5 -.3\n  Rotate -70 1 0 0\n'
          '  Shape "sphere" "float radius" [.5]\nAttributeEnd\n')
    s += ('AttributeBegin\n  Material "glass" "float index" [1.5]\n  Translate .1 -.7 1.2\n'
          '  Shape "sphere" "float radius" [.3]\nAttributeEnd\n')
    return s + "WorldEnd\n"


def manylights(opts, out):
    """An emissive mesh: every triangle of an area-light shape is its own DiffuseAreaLight
    (src/core/api.cpp:1377-1386), so a 24 x 24 grid puts 1152 lights into the spatial light
    distribution -- the device's per-voxel tables switch to their sparse form."""
    name = os.path.splitext(os.path.basename(out))[0]
    cam = 'LookAt 0 1.0 3.6  0 -.2 0  0 1 0\nCamera "perspective" "float fov" [42]'
    s = header(opts, cam, name) + "WorldBegin\n"
    n = int(opts.get("cells", 24))
    P, idx = [], []
    for j in range(n + 1):
        for i in range(n + 1):
            x, z = -0.6 + 1.2 * i / n, -0.6 + 1.2 * j / n
            P += [x, 1.5 + 0.05 * math.sin(7 * x) * math.cos(5 * z), z]
    for j in range(n):
        for i in range(n):
            a = j * (n + 1) + i
            idx += [a, a + 1, a + n + 2, a, a + n + 2, a + n + 1]
    fmt = lambda v: " ".join(f"{x:.6g}" for x in v)
    s += ('AttributeBegin\n  AreaLightSource "diffuse" "blackbody L" [5000 12]\n'
          f'  Shape "trianglemesh" "integer indices" [{" ".join(map(str, idx))}]\n    "point P" [{fmt(P)}]\nAttributeEnd\n')
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {WHITE}\n'
    s += "  " + quad((-3, -1, 3), (3, -1, 3), (3, -1, -3), (-3, -1, -3)) + "AttributeEnd\n"
    s += (f'AttributeBegin\n  Material "matte" "spectrum Kd" {RED}\n  Translate -.8 -.55 0\n'
          '  Shape "sphere" "float radius" [.45]\nAttributeEnd\n')
    s += ('AttributeBegin\n  Material "plastic" "spectrum Kd" [400 .2 700 .6] "float roughness" [.1]\n  Translate .7 -.6 .3\n'
          '  Shape "sphere" "float radius" [.4]\nAttributeEnd\n')
    return s + "WorldEnd\n"


def cornell(opts, out):
    cam = 'LookAt 0 0 3.8  0 0 0  0 1 0\nCamera "perspective" "float fov" [39]'
    return header(opts, cam, os.path.splitext(os.path.basename(out))[0]) + cornell_world()


def cornell_lens(opts, out):
    lens = os.path.join(os.path.dirname(os.path.abspath(out)), "dgauss.50mm.dat")
    with open(lens, "w") as f:
        f.write(DGAUSS_50MM)
    cam = ('LookAt 0 0 3.8  0 0 0  0 1 0\nCamera "realistic" "string lensfile" "dgauss.50mm.dat" '
           '"float aperturediameter" [8] "float focusdistance" [3.8] "bool simpleweighting" "true"')
    if opts.get("ca") == "true":  # wavelength-dependent lens refraction (cameras/realistic.cpp:352-358)
        cam += ' "bool chromaticAberrationEnabled" "true"'
    return header(opts, cam, os.path.splitext(os.path.basename(out))[0]) + cornell_world()


def instanced(opts, out):
    n, k = opts["grid"], opts["inst"]
    xs = np.linspace(-0.5, 0.5, n)
    X, Y = np.meshgrid(xs, xs, indexing="xy")
    Z = 0.08 * np.sin(9 * X) * np.cos(7 * Y) + 0.03 * np.sin(31 * X * Y)
    P = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    idx = []
    ii = np.arange(n - 1)
    a = (ii[None, :] + n * ii[:, None]).reshape(-1)
    tri = np.stack([a, a + 1, a + n + 1, a, a + n + 1, a + n], -1).reshape(-1)
    cam = f'LookAt 0 {-0.9 * k} {0.55 * k}  0 0 0  0 0 1\nCamera "perspective" "float fov" [45]'
    s = header(opts, cam, os.path.splitext(os.path.basename(out))[0]) + "WorldBegin\n"
    s += 'LightSource "point" "point from" [0 0 %g] "blackbody I" [6500 %g]\n' % (1.5 * k, 40.0 * k * k)
    s += 'Material "matte" "spectrum Kd" [400 .6 700 .5]\n'
    s += 'ObjectBegin "tile"\n  Shape "trianglemesh" "integer indices" [\n'
    s += " ".join(str(int(v)) for v in tri) + ' ]\n  "point P" [\n'
    s += " ".join(f"{v:.6g}" for v in P.reshape(-1)) + " ]\nObjectEnd\n"
    rng = np.random.RandomState(7)
    for j in range(k):
        for i in range(k):
            ang = rng.uniform(-8, 8)
            s += (f"AttributeBegin\n  Translate {i - (k - 1) / 2:g} {j - (k - 1) / 2:g} "
                  f"{rng.uniform(-.05, .05):.4f}\n  Rotate {ang:.3f} 0 0 1\n"
                  f'  ObjectInstance "tile"\nAttributeEnd\n')
    s += "WorldEnd\n"
    return s


def zoo(opts, out):
    """One of everything the device path supports, for parity: thin-lens perspective camera,
    crop window, Oren-Nayar / mirror / rough plastic / anisotropic metal / glass, mesh with
    normals + uv + tangents, partial sphere, sphere and two-sided quad area lights, a point
    light, a reversed-orientation instance, power light sampling, optional wide filter."""
    w, h = opts["res"]
    name = os.path.splitext(os.path.basename(out))[0]
    s = 'LookAt 0 1.2 4.2  0 0.2 0  0 1 0\n'
    s += 'Camera "perspective" "float fov" [40] "float lensradius" [0.03] "float focaldistance" [4.2]\n'
    if opts.get("filter", "box") == "gaussian":
        s += 'PixelFilter "gaussian" "float xwidth" [1.5] "float ywidth" [1.5]\n'
    s += (f'Film "image" "integer xresolution" [{w}] "integer yresolution" [{h}] '
          f'"float cropwindow" [0.1 0.95 0.05 0.9] "string filename" "{name}.exr"\n')
    s += f'Sampler "halton" "integer pixelsamples" [{opts["spp"]}]\n'
    s += (f'Integrator "path" "integer maxdepth" [{opts["maxdepth"]}] '
          f'"string lightsamplestrategy" "{opts["strategy"]}"\n')
    s += "WorldBegin\n"
    s += 'LightSource "point" "point from" [-1.5 2.5 1.5] "blackbody I" [4000 6]\n'
    # sphere area light
    s += ('AttributeBegin\n  AreaLightSource "diffuse" "blackbody L" [6500 8]\n  Translate 1.2 1.8 -0.5\n'
          '  Shape "sphere" "float radius" [0.25]\nAttributeEnd\n')
    # two-sided quad light
    s += ('AttributeBegin\n  AreaLightSource "diffuse" "spectrum L" [400 2 550 4 700 6] "bool twosided" "true"\n  '
          + quad((-0.4, 2.2, -1.2), (0.4, 2.2, -1.2), (0.4, 1.7, -1.0), (-0.4, 1.7, -1.0)) + "AttributeEnd\n")
    # floor: Oren-Nayar, mesh with uv and normals
    s += ('AttributeBegin\n  Material "matte" "spectrum Kd" [400 .3 700 .7] "float sigma" [25]\n'
          '  Shape "trianglemesh" "integer indices" [0 1 2 0 2 3] "point P" [-3 0 -3 3 0 -3 3 0 3 -3 0 3]\n'
          '    "float uv" [0 0 1 0 1 1 0 1] "normal N" [0 1 0 0 1 0 0 1 0 0 1 0]\nAttributeEnd\n')
    # back wall: plain matte
    s += f'AttributeBegin\n  Material "matte" "spectrum Kd" {GREEN}\n  ' + \
        quad((-3, 0, -2), (3, 0, -2), (3, 3, -2), (-3, 3, -2)) + "AttributeEnd\n"
    # mirror sphere
    s += ('AttributeBegin\n  Material "mirror" "spectrum Kr" [400 .9 700 .8]\n  Translate -1.3 0.45 0\n'
          '  Shape "sphere" "float radius" [0.45]\nAttributeEnd\n')
    # rough plastic sphere
    s += ('AttributeBegin\n  Material "plastic" "spectrum Kd" [400 .1 700 .6] "spectrum Ks" [400 .4 700 .4] '
          '"float roughness" [0.15]\n  Translate -0.3 0.4 0.6\n  Shape "sphere" "float radius" [0.4]\nAttributeEnd\n')
    # glass partial sphere (clipped in z and phi), rotated
    s += ('AttributeBegin\n  Material "glass" "float index" [1.45]\n  Translate 0.7 0.5 0.2\n  Rotate 35 1 0 0\n'
          '  Shape "sphere" "float radius" [0.5] "float zmin" [-0.3] "float zmax" [0.4] "float phimax" [300]\n'
          'AttributeEnd\n')
    # anisotropic metal on a curved mesh with normals, uv and tangents
    n = 9
    us = np.linspace(0, 1, n)
    P, N, S, UV, idx = [], [], [], [], []
    for j in range(n):
        for i in range(n):
            a = (us[i] - 0.5) * 1.6
            P += [1.9 + 0.45 * math.sin(a), 0.15 + us[j] * 0.9, -0.4 + 0.45 * math.cos(a) - 0.45]
            N += [math.sin(a), 0, math.cos(a)]
            S += [math.cos(a), 0, -math.sin(a)]
            UV += [us[i], us[j]]
    for j in range(n - 1):
        for i in range(n - 1):
            a = j * n + i
            idx += [a, a + 1, a + n + 1, a, a + n + 1, a + n]
    fmt = lambda v: " ".join(f"{x:.6g}" for x in v)
    s += ('AttributeBegin\n  Material "metal" "float uroughness" [0.05] "float vroughness" [0.25]\n'
          f'  Shape "trianglemesh" "integer indices" [{" ".join(map(str, idx))}]\n    "point P" [{fmt(P)}]\n'
          f'    "normal N" [{fmt(N)}]\n    "vector S" [{fmt(S)}]\n    "float uv" [{fmt(UV)}]\nAttributeEnd\n')
    # an instanced, reversed-orientation object
    s += ('ObjectBegin "blob"\n  Material "matte" "spectrum Kd" ' + RED + '\n  ReverseOrientation\n'
          '  Shape "trianglemesh" "integer indices" [0 1 2 0 2 3 0 3 1 1 3 2]\n'
          '    "point P" [0 0 0 .4 0 0 .2 0 .35 .2 .5 .12]\nObjectEnd\n')
    for k, (x, z, r) in enumerate([(-2.2, 0.8, 20), (2.3, 1.0, -50)]):
        s += f'AttributeBegin\n  Translate {x} 0.001 {z}\n  Rotate {r} 0 1 0\n  Scale 1.5 1.5 1.5\n  ObjectInstance "blob"\nAttributeEnd\n'
    s += "WorldEnd\n"
    return s


def furnace(opts, out, area):
    # src/tests/analytic_scenes.cpp:71-98 (point light, Kd = 0.5) and :135-165
    # (Kd = 0.5 sphere that also emits Le = 0.5): a camera inside a unit
    # sphere; every pixel's converged radiance is 1.
    cam = 'LookAt 0 0 0  0 0 1  0 1 0\nCamera "perspective" "float fov" [45]'
    s = header(opts, cam, os.path.splitext(os.path.basename(out))[0]) + "WorldBegin\n"
    if area:
        s += 'AreaLightSource "diffuse" "spectrum L" [400 .5 700 .5]\n'
    else:
        s += f'LightSource "point" "point from" [0 0 0] "spectrum I" [400 {math.pi:.9g} 700 {math.pi:.9g}]\n'
    s += 'Material "matte" "spectrum Kd" [400 .5 700 .5]\n'
    s += 'ReverseOrientation\nShape "sphere" "float radius" [1]\nWorldEnd\n'
    return s


def main(argv):
    if len(argv) < 3:
        print(__doc__)
        return 2
    name, out = argv[1], argv[2]
    opts = {"res": (64, 64), "spp": 16, "maxdepth": 5, "strategy": "spatial", "grid": 64, "inst": 4}
    for kv in argv[3:]:
        k, v = kv.split("=", 1)
        if k == "res":
            opts[k] = parse_res(v)
        elif k in ("spp", "maxdepth", "grid", "inst"):
            opts[k] = int(v)
        else:
            opts[k] = v
    if name == "cornell":
        text = cornell(opts, out)
    elif name == "lobes_box":
        text = lobes_box(opts, out)
    elif name == "uber_box":
        text = uber_box(opts, out)
    elif name == "cornell_lens":
        text = cornell_lens(opts, out)
    elif name == "instanced":
        text = instanced(opts, out)
    elif name == "zoo":
        text = zoo(opts, out)
    elif name == "envbox":
        text = envbox(opts, out)
    elif name == "texbox":
        text = texbox(opts, out)
    elif name == "manylights":
        text = manylights(opts, out)
    elif name == "furnace_point":
        text = furnace(opts, out, False)
    elif name == "furnace_area":
        text = furnace(opts, out, True)
    else:
        print("unknown scene", name)
        return 2
    with open(out, "w") as f:
        f.write(text)
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))
