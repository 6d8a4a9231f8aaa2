// physics_shim.js -- the one method of the reference whose body changes.
//
// installGpuIntegrate(physics, CONFIG, addon, opts) replaces
// PhysicsSystem.integrate (/root/reference/index.html:247-262) on an existing
// `physics` object.  Everything else in the page -- Body, addBody,
// resolveCollisions, VisualSystem, animate() -- stays as it is: they keep
// reading and writing physics.bodies[*].pos/vel/acc/temperature.
//
//   gravity-only mode (opts.collisions === false): ONE addon.step() per call;
//   default mode: kick1Drift + force on the GPU, the page's OWN
//     resolveCollisions(dt) (idx:256) on the host, kick2 on the GPU -- the
//     order of idx:250-261.
//   opts.collisions === 'gpu-broad-phase': the O(N^2) overlap search of
//     resolveCollisions (idx:301-313) also runs on the GPU (addon.contacts).
//     When no pair of spheres overlaps the host stage is skipped altogether
//     (the reference loop would only empty contactPairs, idx:376-380) and the
//     bodies never leave the device between force and kick-2; when some do, the
//     page's own resolveCollisions(dt) runs unchanged.  A host that wants the
//     response restricted to the candidate pairs can follow the Python mirror
//     (n-body-sim_b200/physics.py, PhysicsSystem.resolveCollisions).
//
// Body objects are marshalled to typed arrays per call (~12 KB at the page's
// N <= 300, idx:57); hosts that own typed arrays already should call the addon
// directly (see INTEGRATION.md).
'use strict';

function installGpuIntegrate(physics, CONFIG, addon, opts = {}) {
  const precision = opts.precision || 'f64';            // the reference computes in fp64
  const Arr = precision === 'f64' ? Float64Array : Float32Array;
  let capacity = opts.capacity || Math.max(CONFIG.maxBodies, 1024);
  let ctx = addon.create(precision, opts.nGpus || 1, capacity);
  let pm = new Arr(4 * capacity), vel = new Arr(4 * capacity), acc = new Arr(4 * capacity), temp = new Arr(capacity);

  let radius = new Arr(capacity);
  const pairScratch = new Uint32Array(2);                // only the count is needed here

  function ensure(n) {
    if (n <= capacity) return;
    addon.destroy(ctx);
    capacity = 2 * n;
    ctx = addon.create(precision, opts.nGpus || 1, capacity);
    pm = new Arr(4 * capacity); vel = new Arr(4 * capacity); acc = new Arr(4 * capacity); temp = new Arr(capacity);
  }

  function pack(bodies, upload = true) {
    const n = bodies.length;
    for (let i = 0; i < n; i++) {
      const b = bodies[i], o = 4 * i;
      pm[o] = b.pos.x; pm[o + 1] = b.pos.y; pm[o + 2] = b.pos.z; pm[o + 3] = b.mass;
      vel[o] = b.vel.x; vel[o + 1] = b.vel.y; vel[o + 2] = b.vel.z;
      acc[o] = b.acc.x; acc[o + 1] = b.acc.y; acc[o + 2] = b.acc.z;
      temp[i] = b.temperature;
    }
    if (upload) addon.upload(ctx, pm.subarray(0, 4 * n), vel.subarray(0, 4 * n), acc.subarray(0, 4 * n), temp.subarray(0, n));
  }

  function unpack(bodies, what, download = true) {
    const n = bodies.length;
    if (download) addon.download(ctx, what.pos ? pm.subarray(0, 4 * n) : null, what.vel ? vel.subarray(0, 4 * n) : null,
                   what.acc ? acc.subarray(0, 4 * n) : null, what.temp ? temp.subarray(0, n) : null);
    for (let i = 0; i < n; i++) {
      const b = bodies[i], o = 4 * i;
      if (what.pos) b.pos.set(pm[o], pm[o + 1], pm[o + 2]);
      if (what.vel) b.vel.set(vel[o], vel[o + 1], vel[o + 2]);
      if (what.acc) b.acc.set(acc[o], acc[o + 1], acc[o + 2]);
      if (what.temp) b.temperature = temp[i];
    }
  }

  const ALL = { pos: true, vel: true, acc: true, temp: true };

  physics.integrate = function integrate(dt) {           // idx:247
    const bodies = this.bodies;
    if (bodies.length === 0) return;
    ensure(bodies.length);
    addon.setHeatDecay(ctx, CONFIG.heatDecay);           // idx:67, read at call time like G and softening
    if (opts.collisions === false) {
      // the whole frame in one addon call: arrays in, integrate(dt), arrays out in place (nbody_step_host)
      const n = bodies.length;
      pack(bodies, false);
      addon.stepHost(ctx, pm.subarray(0, 4 * n), vel.subarray(0, 4 * n), acc.subarray(0, 4 * n), temp.subarray(0, n),
                     dt, CONFIG.G, CONFIG.softening, 1);   // idx:250-261 minus idx:256
      unpack(bodies, ALL, false);
      return;
    }
    pack(bodies);
    addon.kick1Drift(ctx, dt);                           // idx:250-253
    if (opts.collisions === 'gpu-broad-phase') {
      const n = bodies.length;
      if (radius.length < n) radius = new Arr(capacity);
      for (let i = 0; i < n; i++) radius[i] = bodies[i].radius;      // idx:221
      addon.setRadius(ctx, radius.subarray(0, n));
      addon.force(ctx, CONFIG.G, CONFIG.softening);      // idx:255, also tracks every body's nearest neighbour
      if (addon.contacts(ctx, 0, pairScratch) === 0) {   // idx:301-313: nothing overlaps
        this.contactPairs.clear();                       // idx:376-380
        addon.kick2(ctx, dt, CONFIG.heatDecay);          // idx:258-261
        unpack(bodies, ALL);
        return;
      }
    } else {
      addon.force(ctx, CONFIG.G, CONFIG.softening);      // idx:255
    }
    unpack(bodies, ALL);
    this.resolveCollisions(dt);                          // idx:256: the page's own code, unchanged
    pack(this.bodies);                                   // collisions may edit pos/vel and the body list
    addon.kick2(ctx, dt, CONFIG.heatDecay);              // idx:258-261
    unpack(this.bodies, { vel: true, temp: true });
  };

  physics.computeGravity = function computeGravity() {   // idx:264-291
    if (this.bodies.length === 0) return;
    ensure(this.bodies.length);
    pack(this.bodies);
    addon.force(ctx, CONFIG.G, CONFIG.softening);
    unpack(this.bodies, { acc: true });
  };

  return { dispose() { addon.destroy(ctx); }, get ctx() { return ctx; } };
}

module.exports = { installGpuIntegrate };
