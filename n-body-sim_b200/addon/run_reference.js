#!/usr/bin/env node
// run_reference.js -- headless Node/V8 harness for the reference's OWN step
// (SURVEY.md 8c).  It READS /path/to/index.html at run time, slices the Body
// and PhysicsSystem class text out of the module script (idx:208-444) and
// evaluates it with a ~30-line THREE.Vector3 stub (three@0.160.0 is a CDN
// import, idx:24, not on disk), a no-op `visuals`, and the CONFIG literal
// (idx:52-75).  Nothing of the reference is copied into this repo.
//
//   node run_reference.js <index.html> <fixture.bin> <nsteps> [--collisions] [--gpu <nbody_napi.node>]
//
// fixture.bin: header 4 x float64 (n, dt, G, softening) then pos_mass[4n],
// vel[4n], acc[4n], temp[n] as float64 (tests/golden/make_fixture_bin.py).
// Writes <fixture.bin>.out in the same layout and prints one JSON line with the
// timing.  NOT runnable in the build image (no Node.js); kept for a maintainer.
'use strict';
const fs = require('fs');

class Vector3 {
  constructor(x = 0, y = 0, z = 0) { this.x = x; this.y = y; this.z = z; }
  clone() { return new Vector3(this.x, this.y, this.z); }
  copy(v) { this.x = v.x; this.y = v.y; this.z = v.z; return this; }
  set(x, y, z) { this.x = x; this.y = y; this.z = z; return this; }
  add(v) { this.x += v.x; this.y += v.y; this.z += v.z; return this; }
  sub(v) { this.x -= v.x; this.y -= v.y; this.z -= v.z; return this; }
  addScaledVector(v, s) { this.x += v.x * s; this.y += v.y * s; this.z += v.z * s; return this; }
  subVectors(a, b) { this.x = a.x - b.x; this.y = a.y - b.y; this.z = a.z - b.z; return this; }
  addVectors(a, b) { this.x = a.x + b.x; this.y = a.y + b.y; this.z = a.z + b.z; return this; }
  multiplyScalar(s) { this.x *= s; this.y *= s; this.z *= s; return this; }
  divideScalar(s) { return this.multiplyScalar(1 / s); }
  dot(v) { return this.x * v.x + this.y * v.y + this.z * v.z; }
  lengthSq() { return this.x * this.x + this.y * this.y + this.z * this.z; }
  length() { return Math.sqrt(this.lengthSq()); }
  normalize() { return this.divideScalar(this.length() || 1); }
  distanceTo(v) { return Math.sqrt(this.distanceToSquared(v)); }
  distanceToSquared(v) { const dx = this.x - v.x, dy = this.y - v.y, dz = this.z - v.z; return dx * dx + dy * dy + dz * dz; }
  lerp(v, a) { this.x += (v.x - this.x) * a; this.y += (v.y - this.y) * a; this.z += (v.z - this.z) * a; return this; }
  cross(v) { const ax = this.x, ay = this.y, az = this.z; this.x = ay * v.z - az * v.y; this.y = az * v.x - ax * v.z; this.z = ax * v.y - ay * v.x; return this; }
}
class Color { constructor(h) { this.hex = h; } }
const THREE = { Vector3, Color };

function sliceBetween(text, startMarker, endMarker) {
  const a = text.indexOf(startMarker), b = text.indexOf(endMarker, a);
  if (a < 0 || b < 0) throw new Error(`cannot find ${startMarker} .. ${endMarker} in the reference`);
  return text.slice(a, b);
}

function main() {
  const [html, fixture, nstepsArg, ...flags] = process.argv.slice(2);
  const nsteps = parseInt(nstepsArg, 10);
  const src = fs.readFileSync(html, 'utf8');
  const configSrc = sliceBetween(src, 'const CONFIG = {', 'const V3');              // idx:52-75
  const classSrc = sliceBetween(src, 'class Body {', '// 4. VISUALS');  // idx:208-444
  const visuals = { triggerFlash() {}, garbageCollect() {}, spawnExplosion() {}, spawnTrail() {} };
  const factory = new Function('THREE', 'visuals', `${configSrc}\nconst V3 = (x,y,z) => new THREE.Vector3(x,y,z);\n${classSrc}\nreturn { CONFIG, Body, PhysicsSystem };`);
  const { CONFIG, Body, PhysicsSystem } = factory(THREE, visuals);

  const buf = fs.readFileSync(fixture);
  const f = new Float64Array(buf.buffer, buf.byteOffset, buf.byteLength / 8);
  const n = f[0], dt = f[1];
  CONFIG.G = f[2]; CONFIG.softening = f[3];
  const pm = f.subarray(4, 4 + 4 * n), vel = f.subarray(4 + 4 * n, 4 + 8 * n), acc = f.subarray(4 + 8 * n, 4 + 12 * n), temp = f.subarray(4 + 12 * n, 4 + 13 * n);
  const physics = new PhysicsSystem();
  for (let i = 0; i < n; i++) {           // bypass addBody: its 300-body cap (idx:240-242) is not part of the step
    const b = new Body(pm[4 * i + 3], new Vector3(pm[4 * i], pm[4 * i + 1], pm[4 * i + 2]), new Vector3(vel[4 * i], vel[4 * i + 1], vel[4 * i + 2]), 'rock', temp[i]);
    b.acc.set(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2]);
    physics.bodies.push(b);
  }
  if (!flags.includes('--collisions')) physics.resolveCollisions = function () {};   // gravity-only parity (SURVEY 8c)
  const g = flags.indexOf('--gpu');
  if (g >= 0) require('./physics_shim.js').installGpuIntegrate(physics, CONFIG, require(flags[g + 1]), { collisions: flags.includes('--collisions') });

  const t0 = process.hrtime.bigint();
  for (let s = 0; s < nsteps; s++) physics.integrate(dt);
  const sec = Number(process.hrtime.bigint() - t0) * 1e-9;

  const out = new Float64Array(4 + 13 * n);
  out.set(f.subarray(0, 4));
  physics.bodies.forEach((b, i) => {
    out.set([b.pos.x, b.pos.y, b.pos.z, b.mass], 4 + 4 * i);
    out.set([b.vel.x, b.vel.y, b.vel.z, 0], 4 + 4 * n + 4 * i);
    out.set([b.acc.x, b.acc.y, b.acc.z, 0], 4 + 8 * n + 4 * i);
    out[4 + 12 * n + i] = b.temperature;
  });
  fs.writeFileSync(fixture + '.out', Buffer.from(out.buffer));
  console.log(JSON.stringify({ impl: g >= 0 ? 'b200-addon' : 'reference-js', n, nsteps, seconds: sec,
    pairs_per_s: n * (n - 1) / 2 * nsteps / sec, interactions_per_s: n * (n - 1) * nsteps / sec, node: process.version }));
}
main();
