#!/usr/bin/env python
"""Diagnostic: how sparse are the coefficient blocks of the bench images?  For every block the highest
non-zero row and column are classed (<=0, <=1, <=3, <=7); a pruned IDCT would have to be chosen warp-uniformly, so
the table that matters is the class of the worst block among 32 consecutive ones ("warp-max").  Uses the oracle's
decode_capture (checker only).  -> profiles/r02_sparsity_classes.txt"""
import sys, numpy as np
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
from oracle import oracle_lib
from tools import testenc as te
o = oracle_lib.load('reference')
for kind in (0,1,2):
    img = te.synth_rgb(1920,1080,kind,1000+kind)
    s = te.encode(img, subsampling="420", restart_interval=16, quality=90)
    out, coefs, planes, seens = o.decode_capture(s, 3)
    print("kind",kind,"bytes",len(s), "bpp", 8*len(s)/(1920*1080))
    for ci,c in enumerate(coefs):
        b = c.reshape(-1,8,8)!=0
        rmax = np.where(b.any(axis=2), np.arange(8)[None,:], -1).max(axis=1)  # highest nonzero row
        cmax = np.where(b.any(axis=1), np.arange(8)[None,:], -1).max(axis=1)
        def cls(m): return np.digitize(m,[1,2,4])  # 0:<=0(DC row only) 1: <=1, 2: <=3, 3: <=7
        R=cls(rmax); Cc=cls(cmax)
        h=np.zeros((4,4),int)
        np.add.at(h,(R,Cc),1)
        print(" comp",ci,"blocks",len(b)); print((h/len(b)*100).round(1))
        # warp-uniform class: groups of 32 consecutive blocks (row-major) -> max
        n=len(b)//32*32
        Rw=R[:n].reshape(-1,32).max(1); Cw=Cc[:n].reshape(-1,32).max(1)
        hw=np.zeros((4,4),int); np.add.at(hw,(Rw,Cw),1)
        print("  warp-max:"); print((hw/len(Rw)*100).round(1))
