#!/usr/bin/env python
"""float32 simulation (numpy, CPU) of the pilot PLL in the reference's operation order (loop/pll.h:64-70, phase_control_loop.h:58-66)
and in the error form pllErrKernel walks (csrc/demod.cu): largest phase difference and how many 512-sample chunks need wrap
and clamp in the chain, for a clean / noisy / absent / out-of-range pilot.  Diagnosis only."""
import numpy as np, math, sys
f32=np.float32
PI=f32(3.1415926535)
TWO_PI=f32(2.0)*PI
def crit(bw):
    d=math.sqrt(2.0)/2.0; den=1.0+2.0*d*bw+bw*bw
    return f32(4*d*bw/den), f32(4*bw*bw/den)
def ref_loop(x, alpha, beta, lo, hi, phase, freq):
    out=np.empty(len(x),np.float32)
    for i in range(len(x)):
        out[i]=phase
        err=f32(x[i]-phase)
        if err>PI: err=f32(err-TWO_PI)
        elif err<=-PI: err=f32(err+TWO_PI)
        freq=f32(freq+f32(beta*err))
        if freq>hi: freq=hi
        elif freq<lo: freq=lo
        phase=f32(phase+f32(freq+f32(alpha*err)))
        while phase>PI: phase=f32(phase-TWO_PI)
        while phase<-PI: phase=f32(phase+TWO_PI)
    return out, phase, freq
def wrap(v):
    if v>PI: return f32(v-TWO_PI)
    if v<=-PI: return f32(v+TWO_PI)
    return v
MAGIC=f32(12582912.0); INV_2PI=f32(1.0/(2.0*3.1415926535))
def wrap_rint(r):
    # the chain's wrap in pllErrKernel: r - 2 pi rint(r / 2 pi) by the magic-number rounding, three plain FP32 instructions
    t=fma(r,INV_2PI,MAGIC); k=f32(t-MAGIC); return fma(k,-TWO_PI,r)
def fma(a,b,c): return f32(np.float64(a)*np.float64(b)+np.float64(c))
def clampP(p):
    if p>PI: return f32(p-TWO_PI)
    if p<-PI: return f32(p+TWO_PI)
    return p
def err_loop(x, alpha, beta, lo, hi, phase, freq, chunk=512):
    out=np.empty(len(x),np.float32)
    c1=f32(f32(1)-alpha); k=f32(f32(f32(1)-alpha)-beta)
    full = not (freq>lo and freq<hi)
    nfull=0
    for base in range(0,len(x),chunk):
        X=x[base:base+chunk]; m=len(X)
        dx=np.zeros(m,np.float32)
        for i in range(m-1): dx[i]=wrap(f32(X[i+1]-X[i]))
        e0=wrap(f32(X[0]-phase))
        def run(linear):
            e=e0; f=freq; E=np.empty(m,np.float32); fLo=f; fHi=f; rMax=f32(0)
            for i in range(m):
                E[i]=e
                if linear:
                    d=f32(dx[i]-f); f=fma(beta,e,f); fLo=min(fLo,f); fHi=max(fHi,f)
                    e=fma(k,e,d); rMax=max(rMax,abs(e))
                else:
                    g=fma(c1,e,dx[i]); fn=fma(beta,e,f); fLo=min(fLo,fn); fHi=max(fHi,fn)
                    fn=min(max(fn,lo),hi); f=fn; r=f32(g-fn); rMax=max(rMax,abs(r)); e=wrap_rint(r)
            return E,e,f,fLo,fHi,rMax
        ok=False
        if not full:
            E,e,f,fLo,fHi,rMax=run(True)
            ok = rMax<PI and abs(e0)<PI and fLo>=lo and fHi<=hi
        if not ok:
            nfull+=1
            E,e,f,fLo,fHi,rMax=run(False)
            full = not (rMax<PI and fLo>lo and fHi<hi)
        for i in range(m): out[base+i]=clampP(f32(X[i]-E[i]))
        eLast=E[m-1]
        pLast=clampP(f32(X[m-1]-eLast))
        phase=clampP(f32(pLast+f32(f+f32(alpha*eLast))))
        freq=f
    return out, phase, freq, nfull
def pdiff(a,b):
    d=np.abs(a.astype(np.float64)-b.astype(np.float64)); d=np.minimum(d,2*np.pi-d); return d
SCENARIOS=(("clean pilot",0.1,1e-4,19000.0),("noisy pilot",0.1,0.03,19000.0),("very noisy",0.1,0.2,19000.0),("no pilot",0.0,0.1,19000.0),
           ("pilot off limits",0.1,1e-3,19600.0),("edge of limits",0.1,1e-2,19245.0))
def scenario(name,amp,noise,fp,n=60000,sr=250e3,seed=1):
    """both walks over one synthetic pilot; returns (largest phase difference, chunks that needed the full walk, chunks, end-state differences)"""
    alpha,beta=crit(25000.0/sr)
    lo=f32(2*np.pi*18750/sr); hi=f32(2*np.pi*19250/sr); f0=f32(2*np.pi*19000/sr)
    rng=np.random.default_rng(seed)
    t=np.arange(n)/sr
    z=amp*np.exp(1j*(2*np.pi*fp*t+0.7))+noise*(rng.standard_normal(n)+1j*rng.standard_normal(n))
    x=np.arctan2(z.imag.astype(np.float32),z.real.astype(np.float32)).astype(np.float32)
    a,pa,fa=ref_loop(x,alpha,beta,lo,hi,f32(0),f0)
    b,pb,fb,nfull=err_loop(x,alpha,beta,lo,hi,f32(0),f0)
    d=pdiff(a,b)
    return d, nfull, (n+511)//512, abs(float(pa)-float(pb)), abs(float(fa)-float(fb))
if __name__ == "__main__":
    for name,amp,noise,fp in SCENARIOS:
        d,nfull,nch,dp,df=scenario(name,amp,noise,fp)
        print("%-18s max phase diff %.3g rad (mean %.3g), at %d; full-walk chunks %d of %d; end state diff %.3g / %.3g" % (name,d.max(),d.mean(),int(d.argmax()),nfull,nch,dp,df))
