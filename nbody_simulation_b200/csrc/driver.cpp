// driver.cpp — headless stand-in for the reference's sim thread (src/main.rs:110-141): World::new(),
// then a fixed number of World::update(STEP_SIZE) calls, printing what the reference prints once per
// second (`ups`, `step`, `Counting{..}`, main.rs:149-156).  The window / pixels / wgpu presentation
// (main.rs:81-99,143-183) is unavailable on this box (no graphics stack) and is left out; `--ppm`
// writes the headless `draw()` frame (main.rs:41-72) to a file instead.
//
//   nbody_driver [--scene default|disc|plummer|galaxy] [--n N] [--steps K] [--theta T] [--dt D]
//                [--tree quad|lbvh|bvhref] [--exact] [--ppm out.ppm]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/nbody_b200.h"

int main(int argc, char **argv) {
  int scene = NB_SCENE_DEFAULT, tree = NB_TREE_QUAD, force = NB_FORCE_FAST;
  size_t n = 0;
  unsigned steps = 100;
  float theta = 50.0f, dt = 0.1f; // THETA, STEP_SIZE (main.rs:34-35)
  std::string ppm;
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    auto next = [&]() -> const char * { return i + 1 < argc ? argv[++i] : ""; };
    if (a == "--scene") {
      std::string s = next();
      scene = s == "disc" ? NB_SCENE_UNIFORM_DISC : s == "plummer" ? NB_SCENE_PLUMMER
              : s == "galaxy" ? NB_SCENE_TWO_GALAXY : NB_SCENE_DEFAULT;
    } else if (a == "--n") n = strtoull(next(), nullptr, 10);
    else if (a == "--steps") steps = (unsigned)strtoul(next(), nullptr, 10);
    else if (a == "--theta") theta = strtof(next(), nullptr);
    else if (a == "--dt") dt = strtof(next(), nullptr);
    else if (a == "--tree") {
      const std::string t = next(); // bvhref = the reference's live tree (mod bvh_tree, main.rs:1), bit for bit
      tree = t == "lbvh" ? NB_TREE_LBVH : t == "bvhref" ? NB_TREE_BVH_REF : NB_TREE_QUAD;
    }
    else if (a == "--exact") force = NB_FORCE_EXACT;
    else if (a == "--ppm") ppm = next();
    else { fprintf(stderr, "unknown argument %s\n", a.c_str()); return 2; }
  }
  size_t total = nb_scene_generate(scene, 1, n, nullptr, 0);
  std::vector<nb_particle> parts(total);
  nb_scene_generate(scene, 1, n, parts.data(), total);
  printf("len: %zu\n", total); // main.rs:343

  nb_config cfg;
  nb_default_config(&cfg, total);
  cfg.theta = theta; cfg.dt = dt; cfg.tree_kind = tree; cfg.force_mode = force;
  nb_ctx *ctx = nb_create(&cfg);
  if (!ctx) { fprintf(stderr, "nb_create: %s\n", nb_last_error(nullptr)); return 1; }
  if (nb_set_bodies_aos(ctx, parts.data(), total) != NB_OK) { fprintf(stderr, "%s\n", nb_last_error(ctx)); return 1; }

  nb_counting counter;
  memset(&counter, 0, sizeof(counter));
  auto t0 = std::chrono::steady_clock::now(), last = t0;
  unsigned long long last_updates = 0;
  for (unsigned done = 0; done < steps;) {
    unsigned batch = steps - done < 16 ? steps - done : 16;
    if (nb_step(ctx, batch, &counter) != NB_OK) { fprintf(stderr, "nb_step: %s\n", nb_last_error(ctx)); return 1; }
    done += batch;
    auto now = std::chrono::steady_clock::now();
    if (std::chrono::duration<double>(now - last).count() >= 1.0 || done == steps) {
      printf("ups: %llu\nstep: %u\nCounting { build_bvh: %.6f, sum_gravity: %.6f, post_calculations: %.6f }\n",
             (unsigned long long)done - last_updates, done, counter.build_s, counter.traverse_s, counter.integrate_s);
      last_updates = done;
      last = now;
    }
  }
  double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  printf("steps/s: %.2f  interactions/s: %.4g  (%u steps, %.3f s wall)\n", steps / wall,
         (double)counter.interactions / wall, steps, wall);
  if (!ppm.empty()) {
    const unsigned R = 1250; // RENDER_HEIGHT, main.rs:32
    std::vector<unsigned char> frame(4ull * R * R);
    if (nb_draw(ctx, frame.data(), R) != NB_OK) { fprintf(stderr, "nb_draw: %s\n", nb_last_error(ctx)); return 1; }
    FILE *f = fopen(ppm.c_str(), "wb");
    if (!f) { perror("ppm"); return 1; }
    fprintf(f, "P6\n%u %u\n255\n", R, R);
    for (size_t p = 0; p < (size_t)R * R; ++p) { // premultiply by alpha over black, as a viewer would
      unsigned a = frame[4 * p + 3];
      unsigned char rgb[3] = {(unsigned char)(frame[4 * p] * a / 255), (unsigned char)(frame[4 * p + 1] * a / 255),
                              (unsigned char)(frame[4 * p + 2] * a / 255)};
      fwrite(rgb, 1, 3, f);
    }
    fclose(f);
  }
  nb_destroy(ctx);
  return 0;
}
