// tpimd5sum -- command-line front end for the VM-side loops of TPI's machine script.
//
// The script (/root/reference/task/common/machine/machine-script.sh.tpl) detects change with
//   md5sum "$TPI_LOG_DIRECTORY"/*                                   (:111, every 5 s)
//   find "$TPI_DATA_DIRECTORY" -printf "%T@\n" | sort | tail -1     (:119, every 10 s)
// and then runs a full `rclone sync`.  This tool answers the same questions from a GPU-built manifest:
//   tpimd5sum FILE...                     the md5sum lines for FILE... (same text as coreutils md5sum)
//   tpimd5sum --manifest DIR [-x RULE]... [-o FILE.tpim]
//                                         md5sum lines of every file machine.Transfer would move from DIR
//                                         (built-in excludes + RULEs), optionally persisted
//   tpimd5sum --changed OLD.tpim DIR [-x RULE]... [-o NEW.tpim]
//                                         which objects a checkpoint sync must upload: "+ path" new,
//                                         "* path" content changed; exit status 1 if any, 0 if none
// All digests are computed on the GPU (device 0, or $TPIMD5_DEVICE); there is no CPU fallback.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <sys/stat.h>
#include <unistd.h>
#include <vector>

#include "../include/tpimd5.h"

static int die(const char *what) {
    fprintf(stderr, "tpimd5sum: %s: %s\n", what, tpimd5_last_error());
    return 2;
}

static void print_line(const uint8_t md5[16], const char *path) {
    char hex[33];
    tpimd5_hex(md5, hex);
    const bool esc = strpbrk(path, "\\\n\r") != nullptr;
    if (esc) fputc('\\', stdout);
    fputs(hex, stdout);
    fputs("  ", stdout);
    for (const char *p = path; *p; ++p) {
        if (esc && *p == '\\') fputs("\\\\", stdout);
        else if (esc && *p == '\n') fputs("\\n", stdout);
        else if (esc && *p == '\r') fputs("\\r", stdout);
        else fputc(*p, stdout);
    }
    fputc('\n', stdout);
}

int main(int argc, char **argv) {
    std::vector<const char *> files, rules;
    const char *mode_dir = nullptr, *old_file = nullptr, *out_file = nullptr;
    bool manifest = false, changed = false;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--manifest" && i + 1 < argc) { manifest = true; mode_dir = argv[++i]; }
        else if (a == "--changed" && i + 2 < argc) { changed = true; old_file = argv[++i]; mode_dir = argv[++i]; }
        else if (a == "-x" && i + 1 < argc) rules.push_back(argv[++i]);
        else if (a == "-o" && i + 1 < argc) out_file = argv[++i];
        else if (a == "-h" || a == "--help") {
            puts("usage: tpimd5sum FILE... | --manifest DIR [-x RULE]... [-o OUT.tpim] | --changed OLD.tpim DIR [-x RULE]... [-o NEW.tpim]");
            return 0;
        } else files.push_back(argv[i]);
    }
    const char *dev_env = getenv("TPIMD5_DEVICE");
    int dev = dev_env ? atoi(dev_env) : 0;
    tpimd5_ctx *ctx = nullptr;
    if (tpimd5_init(&dev, 1, 64u << 20, &ctx) != 0) return die("init");
    int status = 0;
    if (!manifest && !changed) {
        if (files.empty()) { fprintf(stderr, "tpimd5sum: no files\n"); tpimd5_destroy(ctx); return 2; }
        std::vector<uint8_t> dig(16 * files.size());
        size_t failed = (size_t)-1;
        if (tpimd5_hash_files(ctx, files.data(), nullptr, files.size(), dig.data(), &failed) != 0) { status = die("hash"); }
        else for (size_t i = 0; i < files.size(); ++i) print_line(&dig[16 * i], files[i]);
    } else {
        tpimd5_filter *flt = nullptr;
        tpimd5_manifest *m = nullptr, *old = nullptr;
        if (tpimd5_filter_new(rules.data(), rules.size(), nullptr, &flt) != 0) status = die("filter");
        else if (tpimd5_manifest_scan(ctx, mode_dir, flt, &m) != 0) status = die("scan");
        else if (manifest) {
            for (size_t i = 0; i < tpimd5_manifest_count(m); ++i) {
                const char *p; uint8_t d[16]; int has = 0;
                tpimd5_manifest_entry(m, i, &p, nullptr, nullptr, d, &has);
                if (has) print_line(d, p);
            }
        } else if (access(old_file, F_OK) != 0 ? tpimd5_manifest_from_arrays(nullptr, nullptr, nullptr, nullptr, nullptr, 0, &old) != 0
                                               : tpimd5_manifest_load(old_file, &old) != 0) {
            status = die("load");      // (no old manifest yet -- the loop's first turn -- counts as an empty one: everything is new)
        } else {
            const size_t n = tpimd5_manifest_count(m);
            std::vector<uint8_t> act(n + 1);
            size_t nt = 0, gone = 0;
            if (tpimd5_plan_transfer(ctx, m, old, 1 /* --checksum */, 1, act.data(), &nt, &gone) != 0) status = die("plan");
            else {
                for (size_t i = 0; i < n; ++i) {
                    if (act[i] != TPIMD5_ACT_TRANSFER_NEW && act[i] != TPIMD5_ACT_TRANSFER) continue;
                    const char *p;
                    tpimd5_manifest_entry(m, i, &p, nullptr, nullptr, nullptr, nullptr);
                    printf("%c %s\n", act[i] == TPIMD5_ACT_TRANSFER_NEW ? '+' : '*', p);
                }
                if (gone) printf("- %zu object(s) only in the old manifest\n", gone);
                status = (nt || gone) ? 1 : 0;
            }
        }
        if (status != 2 && out_file && m && tpimd5_manifest_save(m, out_file) != 0) status = die("save");
        tpimd5_manifest_free(old);
        tpimd5_manifest_free(m);
        tpimd5_filter_free(flt);
    }
    tpimd5_destroy(ctx);
    return status;
}
