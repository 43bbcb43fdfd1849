// ref_golden.rs — run INSIDE the reference crate (KristinnVikarJ/nbody-simulation) to pin this repo's CPU oracle
// against the real Rust code.  It cannot be compiled in the build container of this repo (no cargo / rustc, the
// 232 crates of Cargo.lock are not vendored); a maintainer with a Rust toolchain runs it once:
//
//   1. python tools/ref_golden/export_particles.py            # writes tools/ref_golden/particles.bin (+ .json)
//   2. append this file to the reference's src/main.rs:        cat ref_golden.rs >> src/main.rs
//      (it must live in main.rs: `mod bvh_tree` and `World` are private to the binary crate)
//   3. NB_GOLDEN_DIR=/path/to/tools/ref_golden cargo test --release ref_golden -- --nocapture
//   4. python tools/ref_golden/to_npz.py                       # -> tools/ref_golden/reference_golden.npz
//   5. python -m pytest tests/test_golden.py --reference-golden tools/ref_golden/reference_golden.npz
//
// What it dumps (little endian, no headers; sizes follow from particles.json):
//   ref_nodes.bin   per node of BVHTree::from(...) + calculate_gravity(), DEPTH-FIRST PRE-ORDER (child 0 first):
//                   u32 depth, u32 is_leaf, u32 start (index of the node's first particle in the permuted Vec),
//                   u32 count, u32 total_mass, f32 cog.x, cog.y, f32 offset.x, offset.y, f32 size.x, size.y
//   ref_vec.bin     the Vec<Particle> AFTER BVHTree::from permuted it: per particle f32 x, y, vx, vy, u32 weight
//   ref_accel.bin   f32 ax, ay per particle of the ORIGINAL (cloned) order, exactly what World::update collects
//   ref_traj.bin    the Vec<Particle> after STEPS calls of World::update(DT, ..) AS SHIPPED (acceleration[idx] is
//                   applied to the permuted Vec: the index mismatch of main.rs:398-423 is part of the reference)
//   ref_meta.txt    THETA as compiled, STEPS, DT, node count
// THETA is a compile-time const (main.rs:35): the dump is valid for the value the crate was built with; set it to
// the `theta` of particles.json (0.5) before step 3 to compare with tests/golden/golden_small.npz.
#[cfg(test)]
mod ref_golden {
    use super::*;
    use crate::bvh_tree::BVHTree;
    use std::io::{Read, Write};

    fn dir() -> String {
        std::env::var("NB_GOLDEN_DIR").unwrap_or_else(|_| ".".to_string())
    }

    fn read_particles() -> Vec<Particle> {
        let mut buf = Vec::new();
        std::fs::File::open(format!("{}/particles.bin", dir()))
            .expect("particles.bin (run export_particles.py first)")
            .read_to_end(&mut buf)
            .unwrap();
        buf.chunks_exact(24)
            .map(|c| {
                let f = |i: usize| f32::from_le_bytes([c[4 * i], c[4 * i + 1], c[4 * i + 2], c[4 * i + 3]]);
                Particle {
                    position: vec2f(f(0), f(1)),
                    velocity: vec2f(f(2), f(3)),
                    weight: u32::from_le_bytes([c[16], c[17], c[18], c[19]]),
                }
            })
            .collect()
    }

    fn put_f32(out: &mut Vec<u8>, v: f32) {
        out.extend_from_slice(&v.to_le_bytes());
    }
    fn put_u32(out: &mut Vec<u8>, v: u32) {
        out.extend_from_slice(&v.to_le_bytes());
    }

    fn dump_vec(path: &str, ps: &[Particle]) {
        let mut out = Vec::new();
        for p in ps {
            put_f32(&mut out, p.position.x());
            put_f32(&mut out, p.position.y());
            put_f32(&mut out, p.velocity.x());
            put_f32(&mut out, p.velocity.y());
            put_u32(&mut out, p.weight);
        }
        std::fs::File::create(path).unwrap().write_all(&out).unwrap();
    }

    // pre-order walk; `base` is the address of the Vec's first particle, so that a leaf's slice gives its start index
    fn dump_tree(t: &BVHTree, depth: u32, base: *const Particle, next_start: &mut u32, out: &mut Vec<u8>) -> u32 {
        match t {
            BVHTree::Leaf { children, boundary } => {
                let start = if children.is_empty() {
                    *next_start
                } else {
                    ((children.as_ptr() as usize - base as usize) / std::mem::size_of::<Particle>()) as u32
                };
                *next_start = start + children.len() as u32;
                put_u32(out, depth);
                put_u32(out, 1);
                put_u32(out, start);
                put_u32(out, children.len() as u32);
                put_u32(out, t.get_total_mass());
                let c = t.get_center_of_gravity();
                put_f32(out, c.x());
                put_f32(out, c.y());
                put_f32(out, boundary.offset.x());
                put_f32(out, boundary.offset.y());
                put_f32(out, boundary.size.x());
                put_f32(out, boundary.size.y());
                1
            }
            BVHTree::Root { boundary, total_mass, children, center_of_gravity } => {
                // start / count are patched by to_npz.py from the leaves below (the Root holds no slice)
                put_u32(out, depth);
                put_u32(out, 0);
                put_u32(out, *next_start);
                put_u32(out, 0);
                put_u32(out, *total_mass);
                put_f32(out, center_of_gravity.x());
                put_f32(out, center_of_gravity.y());
                put_f32(out, boundary.offset.x());
                put_f32(out, boundary.offset.y());
                put_f32(out, boundary.size.x());
                put_f32(out, boundary.size.y());
                let a = dump_tree(&children[0], depth + 1, base, next_start, out);
                let b = dump_tree(&children[1], depth + 1, base, next_start, out);
                1 + a + b
            }
        }
    }

    #[test]
    fn ref_golden() {
        const STEPS: usize = 3;
        const DT: f32 = 0.1;
        let original = read_particles();

        // (1) tree + permuted Vec + forces of ONE build, as World::update does it (main.rs:398-416)
        let mut vec = original.clone();
        let cloned = vec.clone();
        let base = vec.as_ptr();
        let mut nodes = Vec::new();
        let n_nodes;
        let mut accel = Vec::new();
        {
            let mut tree = BVHTree::from(vec.as_mut());
            tree.calculate_gravity();
            let mut next = 0u32;
            n_nodes = dump_tree(&tree, 0, base, &mut next, &mut nodes);
            for p in cloned.iter() {
                let mut a = vec2f(0.0, 0.0);
                World::bvh_sum_gravity(&p.clone().into(), &tree, &mut a);
                put_f32(&mut accel, a.x());
                put_f32(&mut accel, a.y());
            }
        }
        std::fs::File::create(format!("{}/ref_nodes.bin", dir())).unwrap().write_all(&nodes).unwrap();
        std::fs::File::create(format!("{}/ref_accel.bin", dir())).unwrap().write_all(&accel).unwrap();
        dump_vec(&format!("{}/ref_vec.bin", dir()), &vec);

        // (2) STEPS x World::update as shipped
        let mut world = World { particles: original.clone() };
        let mut counter = Counting { build_bvh: 0.0, sum_gravity: 0.0, post_calculations: 0.0 };
        for _ in 0..STEPS {
            world.update(DT, &mut counter);
        }
        dump_vec(&format!("{}/ref_traj.bin", dir()), &world.particles);

        let meta = format!("theta {}\nsteps {}\ndt {}\nnodes {}\nparticles {}\n", THETA, STEPS, DT, n_nodes, original.len());
        std::fs::File::create(format!("{}/ref_meta.txt", dir())).unwrap().write_all(meta.as_bytes()).unwrap();
        println!("ref_golden: {} particles, {} nodes, THETA = {}", original.len(), n_nodes, THETA);
    }
}
