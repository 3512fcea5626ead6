"""Generates the committed golden fixtures under tests/golden/ from the
reference's OWN test files.  Run in the dev container only (reads
/root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

Sources (reference file:line):
  * tests/test_utils/test_overlaps.py:7-14      -> overlaps_golden.json
  * tests/test_data/test_datasets/test_dota.py:13-21,55 and
    tests/data/labelTxt/P0004__1__0___0.txt     -> dota_p0004_golden.json
    (gt polygons -> 'oc' OBBs exactly as mmrotate/core/bbox/transforms.py:334-357
    `poly2obb_np_oc` does: cv2.minAreaRect + angle normalisation; loader dtype
    float32 as at mmrotate/datasets/dota.py:100-118)
  * tests/test_data/test_pipelines/test_rtransforms.py:140-158 -> iof_golden.json
"""
import ast
import json
import os
import re

import cv2
import numpy as np

REF = '/root/reference'
HERE = os.path.dirname(os.path.abspath(__file__))


def _literal(src, name):
    for node in ast.parse(src).body:
        if isinstance(node, ast.Assign) and getattr(node.targets[0], 'id', None) == name:
            return ast.literal_eval(node.value)
    raise KeyError(name)


def overlaps():
    src = open(os.path.join(REF, 'tests/test_utils/test_overlaps.py')).read()
    out = dict(predict=_literal(src, 'predict'), gt=_literal(src, 'gt'),
               expect_ious=_literal(src, 'expect_ious'), atol=1e-3,
               source='tests/test_utils/test_overlaps.py:7-14')
    json.dump(out, open(os.path.join(HERE, 'overlaps_golden.json'), 'w'), indent=1)


def poly2obb_oc(poly):
    pts = np.array(poly, dtype=np.float32).reshape(4, 2)
    (x, y), (w, h), a = cv2.minAreaRect(pts)
    if w < 2 or h < 2:
        return None
    while not 0 < a <= 90:
        if a == -90:
            a += 180
        else:
            a += 90
            w, h = h, w
    a = a / 180 * np.pi
    return [x, y, w, h, a]


def dota():
    gts, polys = [], []
    for line in open(os.path.join(REF, 'tests/data/labelTxt/P0004__1__0___0.txt')):
        f = line.split()
        poly = [float(v) for v in f[:8]]
        obb = poly2obb_oc(poly)
        if obb is None:
            continue
        polys.append(poly)
        gts.append([float(np.float32(v)) for v in obb])
    src = open(os.path.join(REF, 'tests/test_data/test_datasets/test_dota.py')).read()
    m = re.search(r'np\.array\((\[\[.*?\]\])\)', src, re.S)
    dets = ast.literal_eval(m.group(1))
    out = dict(gt_polys=polys, gt_obbs_oc_f32=gts, dets=dets, expected_mAP=0.7272727,
               decimals=7, cv2_version=cv2.__version__,
               source='tests/test_data/test_datasets/test_dota.py:13-21,55 + '
                      'tests/data/labelTxt/P0004__1__0___0.txt')
    json.dump(out, open(os.path.join(HERE, 'dota_p0004_golden.json'), 'w'), indent=1)


if __name__ == '__main__':
    overlaps()
    dota()
    print('golden fixtures written to', HERE)
