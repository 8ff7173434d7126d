// host/main_headless.cpp -- the reference's main loop (src/main.cpp:173-232) written against the C-ABI of
// include/rc.h, headless: no window, the frame is read back (rc_read_image) or written as PPM.
//
// The reference's own host needs GLFW / GLEW / glm / <windows.h> (build.bat:4-17) and cannot be built in this
// image, so the kept host code (world generation, block store, picking ray) comes from libvoxelhost.so
// (include/voxel_host.h), a from-scratch stand-in that the tests hold to the reference's own code voxel for
// voxel.  INTEGRATION.md shows the same loop as a patch to the real src/main.cpp.
//
//   rc_headless [--frames N] [--size WxH] [--ao S] [--out prefix] [--edit] [--gpu-pick] [--gen-device] [--device D] [--gpus N]
//               [--devices a,b,c] [--still] [--dump-last FILE]
//
// --gpus N: rc_init_multi over devices D .. D+N-1 -- the same loop, one process, one thread; the library deals the bands of
// every frame to its devices and rc_read_image lands each device's bands in the (page-locked) host frame.
// --still keeps the default camera and draws no crosshair (the frame of BASELINE configs 3 / 5); --dump-last FILE writes the last frame raw
// (RGBA8, bottom row first) for byte comparisons.
//
// Per frame: camera yaw advances as rotateCamera() would (src/input.cpp:19-30), the picking ray feeds
// lookingAtBlock (src/main.cpp:81,89), --edit clicks once per 8 frames (src/input.cpp:141-160: delete the picked
// voxel / place one at lastPos) and pushes the change with rc_update_world(RC_ALL_BLOCKS).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <string>
#include <vector>

#include "rc.h"
#include "voxel_host.h"

static void check(int st, const char* what) {  // replaces checkGlError (src/globals.hpp:32-43): report, then stop
    if (st != RC_OK) {
        fprintf(stderr, "%s: %s\n", what, rc_last_error());
        exit(1);
    }
}

// rotateCamera's result for (rotationX, rotationY) degrees (src/input.cpp:25-28), glm's formulas in fp32
static void cameraDirFor(double rotationX, double rotationY, float out[3]) {
    auto normalize = [](float v[3]) {
        const float inv = 1.0f / sqrtf((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]);
        v[0] *= inv, v[1] *= inv, v[2] *= inv;
    };
    const float p = (float)(rotationY * 0.017453292519943295), y = (float)(rotationX * 0.017453292519943295);
    float v[3] = {0.0f, 0.0f * cosf(p) - 1.0f * sinf(p), 0.0f * sinf(p) + 1.0f * cosf(p)};  // rotateX(0,0,1)
    normalize(v);
    float w[3] = {v[0] * cosf(y) + v[2] * sinf(y), v[1], -v[0] * sinf(y) + v[2] * cosf(y)};  // rotateY
    normalize(w);
    memcpy(out, w, sizeof(w));
}

int main(int argc, char** argv) {
    int frames = 16, width = 1920, height = 1080, ao = 64, device = 0, gpus = 1;
    bool edit = false, gpuPick = false, genDevice = false, still = false;
    std::string out, dumpLast, deviceList;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--frames") frames = atoi(next());
        else if (a == "--size") sscanf(next(), "%dx%d", &width, &height);
        else if (a == "--ao") ao = atoi(next());
        else if (a == "--out") out = next();
        else if (a == "--device") device = atoi(next());
        else if (a == "--gpus") gpus = atoi(next());
        else if (a == "--devices") deviceList = next();  // e.g. 0,2,3 (instead of --device D --gpus N)
        else if (a == "--still") still = true;
        else if (a == "--dump-last") dumpLast = next();
        else if (a == "--edit") edit = true;
        else if (a == "--gpu-pick") gpuPick = true;
        else if (a == "--gen-device") genDevice = gpuPick = true;  // no host tree: the picking ray runs on the device too
        else {
            fprintf(stderr, "usage: %s [--frames N] [--size WxH] [--ao S] [--out prefix] [--edit] [--gpu-pick] [--gen-device] [--device D] "
                            "[--gpus N] [--still] [--dump-last file]\n", argv[0]);
            return 2;
        }
    }

    if (gpus > 1 || !deviceList.empty()) {  // createWindow(); setupOpenGl();         src/main.cpp:176-179
        std::vector<int> devices;
        if (deviceList.empty()) {
            for (int i = 0; i < gpus; i++) devices.push_back(device + i);
        } else {
            for (const char* p = deviceList.c_str(); *p;) {
                devices.push_back(atoi(p));
                while (*p && *p != ',') p++;
                if (*p == ',') p++;
            }
            gpus = (int)devices.size();
        }
        check(rc_init_multi(gpus, devices.data()), "rc_init_multi");
    } else {
        check(rc_init(device), "rc_init");
    }
    vh_world* world = vh_create(6);     // initTetraHexaTree();                      :181
    const int64_t seeds[3] = {42, 64, 100};
    bool hostTree = !genDevice;
    auto buildHostTree = [&]() {
        vh_add_startup_blocks(world);
        auto t0 = std::chrono::steady_clock::now();
        vh_generate_terrain(world, 200, 200, seeds);  // genWorld();                 :188-190
        printf("generated chunk! time: %lf\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
        hostTree = true;
    };
    if (genDevice) {  // the same world built on the device, no host tree and no upload
        rc_gen_params gp = {200, 200, 6, 1, {seeds[0], seeds[1], seeds[2]}};
        rc_gen_info gi;
        check(rc_generate_world(&gp, &gi), "rc_generate_world");
        printf("generated chunk on the device! time: %lf (noise %.3f ms + tree %.3f ms on the device; %llu bricks, %llu voxels)\n",
               gi.total_ms / 1e3, gi.heights_ms, gi.build_ms, (unsigned long long)gi.bricks, (unsigned long long)gi.voxels);
    } else {
        buildHostTree();
    }

    float cameraPos[3], cameraDir[3], sun[3];
    vh_default_camera(cameraPos, cameraDir, sun);  // src/globals.cpp:20-23 + normalize :195

    rc_world_desc desc;
    rc_update_info ui;
    if (hostTree) {
        vh_describe(world, &desc);
        check(rc_update_world(&desc, nullptr, 0, nullptr, 0, &ui), "rc_update_world");  // first updateSsboData(): full upload
    }
    rc_world_info wi;
    check(rc_world_info_get(&wi), "rc_world_info_get");
    printf("world: %llu reference nodes (%.1f MB) -> %.2f MB packed, %s %.1f ms, upload %.2f ms\n", (unsigned long long)wi.ref_nodes,
           wi.ref_bytes / 1e6, wi.packed_bytes / 1e6, hostTree ? "flatten" : "generate", wi.flatten_ms, wi.upload_ms);

    // the frame in host memory, page-locked so that every device's read-back runs as an asynchronous copy of its own
    std::vector<uint8_t> frame((size_t)width * height * 4);
    check(rc_host_register(frame.data(), frame.size()), "rc_host_register");
    double rotationX = 45.0, rotationY = 0.0, sumMs = 0.0;  // cameraDir (1,0,1) = 45 degrees of yaw
    const vh_block placed = {vh_rgb(255, 40, 40), 0u, 0.0f};
    for (int f = 0; f < frames; f++) {
        auto start = std::chrono::steady_clock::now();
        // picking ray -> lookingAtBlock                                              src/main.cpp:81
        rc_pick_result pick;
        if (gpuPick) {
            check(rc_pick(cameraPos, cameraDir, 30, &pick), "rc_pick");
        } else {
            int32_t r[7];
            vh_pick(world, cameraPos, cameraDir, 30, r);
            memcpy(pick.pos, r, 12), memcpy(pick.last_pos, r + 3, 12), pick.steps = r[6];
        }
        if (edit && f % 8 == 4 && pick.steps) {  // one click                        src/input.cpp:141-160
            if (!hostTree) buildHostTree();  // the first edit needs the host's tree: built late, diffed against nothing
            if ((f / 8) % 2 == 0) vh_put_block(world, pick.last_pos[0], pick.last_pos[1], pick.last_pos[2], &placed, 6);
            else vh_put_block(world, pick.pos[0], pick.pos[1], pick.pos[2], &placed, 6);
            vh_describe(world, &desc);
            check(rc_update_world(&desc, nullptr, RC_ALL_BLOCKS, nullptr, RC_ALL_BLOCKS, &ui), "rc_update_world");
            printf("\nclick at frame %d: %u nodes / %u arrays changed, %u runs re-emitted, %llu bytes uploaded, %.2f + %.2f ms\n", f,
                   ui.changed_ref_nodes, ui.changed_ref_arrays, ui.reflattened, (unsigned long long)ui.uploaded_bytes, ui.host_ms,
                   ui.upload_ms);
        }
        rc_frame_params p;  // the glUniform* block of render()                      src/main.cpp:83-95
        memset(&p, 0, sizeof(p));
        memcpy(p.origin, cameraPos, 12);
        memcpy(p.camera_dir, cameraDir, 12);
        memcpy(p.sun_dir, sun, 12);
        p.mouse_pos[0] = width * 0.5f, p.mouse_pos[1] = height * 0.5f;  // (--still: no crosshair, the frame of the BASELINE configs)
        memcpy(p.looking_at, pick.pos, 12);
        p.delta_time = (float)(f / 60.0);
        p.width = (uint32_t)width, p.height = (uint32_t)height;
        p.proj_plane[0] = (float)tan(45.0 * 0.017453292519943295);
        p.proj_plane[1] = (float)(tan(45.0 * 0.017453292519943295) * ((float)height) / width);
        p.max_steps = 300, p.shadow_steps = 75, p.ao_samples = (uint32_t)ao, p.ao_steps = 5, p.ao_strength = 1.0f;
        p.flags = (still ? 0u : RC_F_CROSSHAIR) | RC_F_SHADOWS;
        check(rc_render_frame(&p), "rc_render_frame");                           // glDrawArrays :105-107
        check(rc_read_image(frame.data(), frame.size(), RC_RGBA8), "rc_read_image");  // glfwSwapBuffers :216
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - start).count();
        if (f) sumMs += ms;
        if (!out.empty()) {
            char name[512];
            snprintf(name, sizeof(name), "%s%04d.ppm", out.c_str(), f);
            check(rc_write_ppm(name), "rc_write_ppm");
        }
        rc_frame_stats st;
        check(rc_frame_stats_get(&st), "rc_frame_stats_get");
        printf("\rframe %d: %.2f ms host wall (%.3f ms on the device: primary %.3f + ao %.3f), %.1f Mrays", f, ms, st.total_ms,
               st.primary_ms, st.ao_ms, (st.primary_rays + st.shadow_rays + st.ao_rays) / 1e6);
        fflush(stdout);
        if (still) continue;
        rotationX += 25 * (1.0 / 60.0) * 0.5 * 8;  // the right-arrow key held for 8 ticks (src/input.cpp:40,53)
        if (rotationX > 180) rotationX -= 360;
        rotationY = 35.0;  // and the down-arrow until the crosshair is on the ground (the picking ray reaches 30 voxels)
        cameraDirFor(rotationX, rotationY, cameraDir);
    }
    if (!dumpLast.empty()) {
        FILE* fp = fopen(dumpLast.c_str(), "wb");
        if (!fp || fwrite(frame.data(), 1, frame.size(), fp) != frame.size()) {
            fprintf(stderr, "cannot write %s\n", dumpLast.c_str());
            return 1;
        }
        fclose(fp);
    }
    check(rc_host_unregister(frame.data()), "rc_host_unregister");
    if (frames > 1) printf("\nrender time: %.2f ms (%d fps) over %d frames at %dx%d, ao %d, %d GPU%s\n", sumMs / (frames - 1),
                           (int)(1000.0 / (sumMs / (frames - 1))), frames - 1, width, height, ao, gpus, gpus > 1 ? "s" : "");
    vh_destroy(world);
    rc_shutdown();  // glfwDestroyWindow / glfwTerminate                              :229-230
    return 0;
}
