// rt_shade.cuh — per-pixel-frame logic of fragment.glsl (main / trace / raytrace / diffuse / lerp) on top
// of the traversal in rt_device.cuh.  Same numeric contract; every function names the shader lines.
#pragma once
#include "rt_device.cuh"

struct SurfaceHit {
    f3 point, normal, color, emissionColor;
    float emission, roughness;
};

// fragment.glsl:146-174 for the winning triangle only (deferred: the shader does this per candidate,
// the inputs of the winner are the same).
RT_DEV SurfaceHit resolveHit(const DeviceScene& sc, const RayHit& h, f3 o, f3 d) {
    const char* sp = reinterpret_cast<const char*>(sc.shade + h.slot);
    float4 s0, s1, s2, s3;
    ld256(sp, s0, s1);
    ld256(sp + 32, s2, s3);
    const float rough = __ldg(reinterpret_cast<const float*>(sp + 64));
    SurfaceHit r;
    r.point = add3(o, scale3(d, h.t));
    const float w = sub(sub(1.0f, h.u), h.v);
    const f3 n0 = mk3(s0.x, s0.y, s0.z), n1 = mk3(s0.w, s1.x, s1.y), n2 = mk3(s1.z, s1.w, s2.x);
    f3 n = normalize3(add3(add3(scale3(n0, w), scale3(n1, h.u)), scale3(n2, h.v)));
    if (length3(n) < 1e-6f) {
        const float4 t0 = __ldg(reinterpret_cast<const float4*>(sc.tris + h.slot));
        const float4 t1 = __ldg(reinterpret_cast<const float4*>(sc.tris + h.slot) + 1);
        n = normalize3(mk3(t0.w, t1.x, t1.y));
    }
    if (dot3(d, n) > 0.0f) n = neg3(n);
    r.normal = n;
    r.color = mk3(s2.y, s2.z, s2.w);
    r.emissionColor = mk3(s3.x, s3.y, s3.z);
    r.emission = s3.w;
    r.roughness = rough;
    return r;
}

// fragment.glsl:220-232 — three draws, x first.
RT_DEV f3 diffuseDir(f3 normal, uint32_t& seed, bool lambertian) {
    const float dx = randomValue(seed);
    const float dy = randomValue(seed);
    const float dz = randomValue(seed);
    f3 dir = mk3(dx, dy, dz);
    if (!lambertian) {
        if (dot3(dir, normal) < 0.0f) dir = scale3(dir, -1.0f);
        return normalize3(dir);
    }
    return normalize3(add3(normal, dir));
}

// fragment.glsl:234-240 with mix(x,y,a) = x*(1-a) + y*a.
RT_DEV f3 lerpDir(f3 diffuse, f3 specular, float n) {
    const float t = sub(1.0f, n);
    const f3 zero = mk3(0.f, 0.f, 0.f);
    const f3 d0 = length3(diffuse) > 0.0f ? normalize3(diffuse) : zero;
    const f3 d1 = length3(specular) > 0.0f ? normalize3(specular) : zero;
    return normalize3(add3(scale3(d0, sub(1.0f, t)), scale3(d1, t)));
}

// reflect(I,N) = I - (2*dot(N,I))*N
RT_DEV f3 reflect3(f3 I, f3 N) { return sub3(I, scale3(N, mul(2.0f, dot3(N, I)))); }

// Position k of the compact launch domain -> pixel coordinate: the k-th coordinate c with
// (c / tileSize) % numTiles == tile (the shader's tile mask, fragment.glsl:382).
RT_DEV int ownedToPixel(int k, int tileSize, int numTiles, int tile) {
    if (numTiles == 1) return k;
    return (k / tileSize) * (numTiles * tileSize) + tile * tileSize + (k % tileSize);
}

// fragment.glsl:388-400: seed, three warm-up draws, camera direction, jitter.
RT_DEV f3 primaryDirection(const PassParams& P, int px, int py, uint32_t& seed) {
    seed = uint32_t(px) * 1973u ^ uint32_t(py) * 9277u ^ uint32_t(P.frameNumber) * 1664525u;
    randomValue(seed);
    randomValue(seed);
    randomValue(seed);
    const float uvx = __fdiv_rn(add(float(px), 0.5f), float(P.width));
    const float uvy = __fdiv_rn(add(float(py), 0.5f), float(P.height));
    const f3 right = ld3(P.camRight), up = ld3(P.camUp), fwd = ld3(P.camForward);
    const float dx = add(P.dirStartX, mul(uvx, P.xStep));
    const float dy = add(P.dirStartY, mul(uvy, P.yStep));
    f3 dir = normalize3(add3(add3(scale3(right, dx), scale3(up, dy)), scale3(fwd, 1.0f)));
    const float j0 = randomValue(seed);   // camRight takes the first draw (:398)
    const float j1 = randomValue(seed);
    dir = add3(dir, scale3(add3(scale3(right, j0), scale3(up, j1)), P.jitterAmount));
    return normalize3(dir);
}

// fragment.glsl:409-416 (+ the frame-sharded SUM mode of rt_render_frames_sum).
RT_DEV float4 accumulate(const PassParams& P, f3 curr, size_t at) {
    f3 out = curr;
    if (P.sumMode == 2) {
        const float4 p = P.prev[at];
        out = add3(mk3(p.x, p.y, p.z), curr);
    } else if (P.sumMode == 0 && P.frameNumber > 0) {
        const float4 p = P.prev[at];
        const float fn = float(P.frameNumber);
        out = div3(add3(scale3(mk3(p.x, p.y, p.z), fn), curr), float(P.frameNumber + 1));
    }
    return make_float4(out.x, out.y, out.z, 1.0f);
}

// The pixel's texel.  In peer tile mode (rt_peer_open + RT_PASS_TO_PEER) a second copy goes straight into the assembling
// rank's target over NVLink — a posted 16-byte store, so the exchange step of tile mode rides on the render kernel.
RT_DEV void storePixel(const PassParams& P, size_t at, float4 v) {
    P.next[at] = v;
    if (P.peerNext != nullptr) P.peerNext[at] = v;
}
