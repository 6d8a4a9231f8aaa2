// physics_shim.js -- the one method of the reference whose body changes.
//
// installGpuIntegrate(physics, CONFIG, addon, opts) replaces
// PhysicsSystem.integrate (/root/reference/index.html:247-262) on an existing
// `physics` object.  Everything else in the page -- Body, addBody,
// resolveCollisions, VisualSystem, animate() -- stays as it is: they keep
// reading and writing physics.bodies[*].pos/vel/acc/temperature.
//
// Body state is BACKED by the addon's typed arrays (SURVEY.md 8f-2): when the body
// list is bound, the x/y/z of every body's pos, vel and acc Vector3 (idx:214-217)
// and its mass and temperature (idx:211, 223) become accessors over one row of
//     pm[4n] = x,y,z,m   vel[4n]   acc[4n]   temp[n]
// -- the Vector3 objects themselves stay the ones the page created, so every
// reference the page or the renderer holds keeps working, and every Vector3 method
// (addScaledVector, sub, dot, ...) reads and writes the arrays through them.  An
// integrate() call is then ONE addon call on those arrays with no per-body JavaScript
// at all.  The list is re-bound only when it changes: a new array (reset idx:745,
// the collision filter idx:383), another length (spawn idx:835, fragments idx:387-389)
// or other end points (the 300-body shift of addBody, idx:240-242).  Bodies that left
// the list get plain numbers back, so a stale reference cannot alias a row.
//
//   gravity-only mode (opts.collisions === false): ONE addon.stepHost() per call;
//   default mode: upload + kick1Drift + force on the GPU, the page's OWN
//     resolveCollisions(dt) (idx:256) on the host -- working on the arrays through
//     the views --, kick2 on the GPU: the order of idx:250-261.
//   opts.collisions === 'gpu-broad-phase': the O(N^2) overlap search of
//     resolveCollisions (idx:301-313) also runs on the GPU (addon.contacts).
//     When no pair of spheres overlaps the host stage is skipped altogether
//     (the reference loop would only empty contactPairs, idx:376-380) and the
//     bodies never leave the device between force and kick-2; when some do, the
//     page's own resolveCollisions(dt) runs unchanged.
'use strict';

function installGpuIntegrate(physics, CONFIG, addon, opts = {}) {
  const precision = opts.precision || 'f64';            // the reference computes in fp64
  const Arr = precision === 'f64' ? Float64Array : Float32Array;
  let capacity = opts.capacity || Math.max(CONFIG.maxBodies, 1024);
  let ctx = addon.create(precision, opts.nGpus || 1, capacity);
  // the arrays the addon works on; S.bound = the bodies whose fields are views of their rows
  const S = { pm: new Arr(4 * capacity), vel: new Arr(4 * capacity), acc: new Arr(4 * capacity), temp: new Arr(capacity),
              radius: new Arr(capacity), list: null, n: 0, first: null, last: null, bound: [], binds: 0 };
  const pairScratch = new Uint32Array(2);                // only the count is needed here
  function viewNumber(obj, key, name, o) {               // obj[key] <-> S[name][o]
    Object.defineProperty(obj, key, { get: () => S[name][o], set: (val) => { S[name][o] = val; }, configurable: true, enumerable: true });
  }

  function plainNumber(obj, key, value) {                // back to an ordinary data property
    Object.defineProperty(obj, key, { value: value, writable: true, configurable: true, enumerable: true });
  }

  function viewVector(v, name, row) {                    // v.x, v.y, v.z <-> S[name][4 row + 0..2]
    viewNumber(v, 'x', name, 4 * row);
    viewNumber(v, 'y', name, 4 * row + 1);
    viewNumber(v, 'z', name, 4 * row + 2);
  }

  function plainVector(v) {
    const x = v.x, y = v.y, z = v.z;
    plainNumber(v, 'x', x);
    plainNumber(v, 'y', y);
    plainNumber(v, 'z', z);
  }

  function release(b) {
    plainVector(b.pos); plainVector(b.vel); plainVector(b.acc);
    plainNumber(b, 'mass', b.mass);
    plainNumber(b, 'temperature', b.temperature);
  }

  // Bind `bodies` to rows 0..n-1.  Values are read through whatever the bodies carry now (views of their OLD rows, or plain
  // numbers for new bodies) into fresh arrays, so rows can move freely; the old views are released first.
  function bind(bodies) {
    const n = bodies.length;
    if (n > capacity) {
      addon.destroy(ctx);
      capacity = 2 * n;
      ctx = addon.create(precision, opts.nGpus || 1, capacity);
    }
    const pm = new Arr(4 * capacity), vel = new Arr(4 * capacity), acc = new Arr(4 * capacity), temp = new Arr(capacity);
    for (let i = 0; i < n; i++) {
      const b = bodies[i], o = 4 * i;
      pm[o] = b.pos.x; pm[o + 1] = b.pos.y; pm[o + 2] = b.pos.z; pm[o + 3] = b.mass;
      vel[o] = b.vel.x; vel[o + 1] = b.vel.y; vel[o + 2] = b.vel.z;
      acc[o] = b.acc.x; acc[o + 1] = b.acc.y; acc[o + 2] = b.acc.z;
      temp[i] = b.temperature;
    }
    for (const b of S.bound) release(b);                 // (a body that stays is bound again below)
    S.pm = pm; S.vel = vel; S.acc = acc; S.temp = temp;
    if (S.radius.length < capacity) S.radius = new Arr(capacity);
    S.bound = [];
    for (let i = 0; i < n; i++) {
      const b = bodies[i];
      viewVector(b.pos, 'pm', i); viewVector(b.vel, 'vel', i); viewVector(b.acc, 'acc', i);
      viewNumber(b, 'mass', 'pm', 4 * i + 3);
      viewNumber(b, 'temperature', 'temp', i);
      S.bound.push(b);
    }
    S.list = bodies; S.n = n; S.first = n > 0 ? bodies[0] : null; S.last = n > 0 ? bodies[n - 1] : null;
    S.binds += 1;
  }

  function sync(bodies) {                                // O(1) unless the list changed
    const n = bodies.length;
    if (bodies !== S.list || n !== S.n || bodies[0] !== S.first || bodies[n - 1] !== S.last) bind(bodies);
  }

  function upload(n) {
    addon.upload(ctx, S.pm.subarray(0, 4 * n), S.vel.subarray(0, 4 * n), S.acc.subarray(0, 4 * n), S.temp.subarray(0, n));
  }

  function download(n, what) {
    addon.download(ctx, what.pos ? S.pm.subarray(0, 4 * n) : null, what.vel ? S.vel.subarray(0, 4 * n) : null,
                   what.acc ? S.acc.subarray(0, 4 * n) : null, what.temp ? S.temp.subarray(0, n) : null);
  }

  const ALL = { pos: true, vel: true, acc: true, temp: true };

  physics.integrate = function integrate(dt) {           // idx:247
    if (this.bodies.length === 0) return;
    sync(this.bodies);
    let n = S.n;
    addon.setHeatDecay(ctx, CONFIG.heatDecay);           // idx:67, read at call time like G and softening
    if (opts.collisions === false) {
      // the whole step in one addon call on the arrays the bodies ARE (nbody_step_host): no per-body work here
      addon.stepHost(ctx, S.pm.subarray(0, 4 * n), S.vel.subarray(0, 4 * n), S.acc.subarray(0, 4 * n), S.temp.subarray(0, n),
                     dt, CONFIG.G, CONFIG.softening, 1);   // idx:250-261 minus idx:256
      return;
    }
    upload(n);
    addon.kick1Drift(ctx, dt);                           // idx:250-253
    if (opts.collisions === 'gpu-broad-phase') {
      for (let i = 0; i < n; i++) S.radius[i] = S.bound[i].radius;   // idx:221 (a constant of each Body: per list, not per field)
      addon.setRadius(ctx, S.radius.subarray(0, n));
      addon.force(ctx, CONFIG.G, CONFIG.softening);      // idx:255, also tracks every body's nearest neighbour
      if (addon.contacts(ctx, 0, pairScratch) === 0) {   // idx:301-313: nothing overlaps
        this.contactPairs.clear();                       // idx:376-380
        addon.kick2(ctx, dt, CONFIG.heatDecay);          // idx:258-261
        download(n, ALL);
        return;
      }
    } else {
      addon.force(ctx, CONFIG.G, CONFIG.softening);      // idx:255
    }
    download(n, ALL);
    this.resolveCollisions(dt);                          // idx:256: the page's own code, unchanged, on the views
    sync(this.bodies);                                   // merges and fractures replace the list (idx:383-389)
    n = S.n;
    if (n === 0) return;
    upload(n);
    addon.kick2(ctx, dt, CONFIG.heatDecay);              // idx:258-261
    download(n, { vel: true, temp: true });
  };

  physics.computeGravity = function computeGravity() {   // idx:264-291
    if (this.bodies.length === 0) return;
    sync(this.bodies);
    upload(S.n);
    addon.force(ctx, CONFIG.G, CONFIG.softening);
    download(S.n, { acc: true });
  };

  return {
    dispose() {
      for (const b of S.bound) release(b);
      S.bound = []; S.list = null; S.n = 0;
      addon.destroy(ctx);
    },
    get ctx() { return ctx; },
    get binds() { return S.binds; },                     // how many times the body list was (re)bound to the arrays
  };
}

module.exports = { installGpuIntegrate };
