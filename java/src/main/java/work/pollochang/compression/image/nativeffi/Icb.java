package work.pollochang.compression.image.nativeffi;

import java.awt.image.BufferedImage;
import java.awt.image.DataBufferByte;
import java.awt.image.Raster;
import java.io.IOException;
import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_FLOAT;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

/**
 * Panama FFM (java.lang.foreign) binding of libimgcomp_b200.so -- include/imgcomp_b200.h.
 *
 * NOT COMPILED IN THE BUILD IMAGE (no JDK there).  Written against the JDK 21 preview API
 * (JEP 442: compile and run with --enable-preview --enable-native-access=ALL-UNNAMED) and
 * unchanged on JDK 22+ (JEP 454, no preview flag).  The same binding, call for call, is what
 * image_compression_b200/api.py does through ctypes and what the test-suite exercises.
 *
 * One handle per JVM replaces Executors.newFixedThreadPool (CompressionBatch.java:64-68): the
 * pool threads stay, each call binds to a free (GPU, stream, arena) slot inside the library.
 * Every thread keeps ONE page-locked staging buffer for pixels and one for the output
 * (icb_alloc_pinned), grown on demand: the decoded raster lives on the Java heap and has to be
 * copied into native memory anyway, so it is copied straight into memory the GPU can DMA from.
 */
public final class Icb {
    private Icb() { }

    public static final int OK = 0, NO_FIT = 1, ERR_INVALID = -1, ERR_OOM = -2, ERR_CUDA = -3,
            ERR_UNSUPPORTED = -4, ERR_CAPACITY = -5, ERR_NO_DEVICE = -6, ERR_TICKET = -7;
    public static final int FMT_RGB8 = 0, FMT_BGR8 = 1, FMT_RGBA8 = 2, FMT_ABGR8 = 3, FMT_GRAY8 = 4,
            FMT_ARGB32 = 5, FMT_XRGB32 = 6;

    private static final Linker LINKER = Linker.nativeLinker();
    private static final SymbolLookup LIB =
            SymbolLookup.libraryLookup(System.getProperty("imgcomp.lib", "libimgcomp_b200.so"), Arena.global());

    private static MethodHandle h(String name, FunctionDescriptor fd) {
        return LINKER.downcallHandle(LIB.find(name).orElseThrow(() -> new UnsatisfiedLinkError(name)), fd);
    }

    // int icb_init(const int *devices, int n_devices, int slots_per_gpu, int max_w, int max_h, icb_handle **out)
    private static final MethodHandle INIT = h("icb_init",
            FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, ADDRESS));
    private static final MethodHandle SHUTDOWN = h("icb_shutdown", FunctionDescriptor.ofVoid(ADDRESS));
    private static final MethodHandle ALLOC_PINNED = h("icb_alloc_pinned", FunctionDescriptor.of(ADDRESS, JAVA_LONG));
    private static final MethodHandle FREE_PINNED = h("icb_free_pinned", FunctionDescriptor.ofVoid(ADDRESS));
    // int icb_jpg_fit(h, px, w, hgt, pitch, fmt, init_q, target, has_cached, cached_q, cached_scale,
    //                 out, out_cap, *out_len, *best_q, *scale, *used_cache)
    private static final MethodHandle JPG_FIT = h("icb_jpg_fit", FunctionDescriptor.of(JAVA_INT,
            ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_FLOAT, JAVA_LONG,
            JAVA_INT, JAVA_FLOAT, JAVA_DOUBLE, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    // int icb_jpg_encode(h, px, w, hgt, pitch, fmt, q, out, out_cap, *out_len)
    private static final MethodHandle JPG_ENCODE = h("icb_jpg_encode", FunctionDescriptor.of(JAVA_INT,
            ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_FLOAT, ADDRESS, JAVA_LONG, ADDRESS));
    // int icb_png_target_dims(w, hgt, min_w, min_h, *dw, *dh, *scale)
    private static final MethodHandle PNG_DIMS = h("icb_png_target_dims", FunctionDescriptor.of(JAVA_INT,
            JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    // void icb_resize_dims(w, hgt, scale, *dw, *dh)
    private static final MethodHandle RESIZE_DIMS = h("icb_resize_dims",
            FunctionDescriptor.ofVoid(JAVA_INT, JAVA_INT, JAVA_DOUBLE, ADDRESS, ADDRESS));
    // int icb_resize_bilinear(h, src, w, hgt, pitch, fmt, dst, dw, dh, dpitch)
    private static final MethodHandle RESIZE = h("icb_resize_bilinear", FunctionDescriptor.of(JAVA_INT,
            ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT));
    // int icb_png_encode(h, px, w, hgt, pitch, fmt, out, out_cap, *out_len)   (section 8 f3)
    private static final MethodHandle PNG_ENCODE = h("icb_png_encode", FunctionDescriptor.of(JAVA_INT,
            ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS));

    // int icb_png_compress(h, src, w, hgt, pitch, fmt, min_w, min_h, out, out_cap, *out_len, *dw, *dh)
    private static final MethodHandle PNG_COMPRESS = h("icb_png_compress", FunctionDescriptor.of(JAVA_INT,
            ADDRESS, ADDRESS, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS,
            ADDRESS, ADDRESS));

    private static volatile MemorySegment handle;

    /** The process-wide handle; created on first use.  No CUDA device -> IllegalStateException: the
     *  library has no CPU path, the caller decides whether to run the stock Java bodies instead. */
    public static MemorySegment handle() {
        MemorySegment hd = handle;
        if (hd != null) return hd;
        synchronized (Icb.class) {
            if (handle == null) {
                try (Arena a = Arena.ofConfined()) {
                    MemorySegment out = a.allocate(ADDRESS);
                    int slots = Math.max(4, Runtime.getRuntime().availableProcessors() / 2);
                    int rc = (int) INIT.invokeExact(MemorySegment.NULL, 0, slots, 8191, 8191, out);
                    if (rc != OK) throw new IllegalStateException("icb_init failed: " + rc);
                    handle = out.get(ADDRESS, 0);
                    Runtime.getRuntime().addShutdownHook(new Thread(Icb::shutdown));
                } catch (RuntimeException | Error e) {
                    throw e;
                } catch (Throwable t) {
                    throw new IllegalStateException(t);
                }
            }
            return handle;
        }
    }

    public static synchronized void shutdown() {
        if (handle != null) {
            try { SHUTDOWN.invokeExact(handle); } catch (Throwable ignored) { }
            handle = null;
        }
    }

    // ---------------------------------------------------------------- pinned staging, one pair per thread
    private static final class Staging {
        MemorySegment px = MemorySegment.NULL, out = MemorySegment.NULL;

        MemorySegment px(long bytes) { return px = grow(px, bytes); }
        MemorySegment out(long bytes) { return out = grow(out, bytes); }

        private static MemorySegment grow(MemorySegment cur, long bytes) {
            if (cur.byteSize() >= bytes) return cur;
            try {
                if (cur.byteSize() > 0) FREE_PINNED.invokeExact(cur);
                MemorySegment p = (MemorySegment) ALLOC_PINNED.invokeExact(bytes);
                if (p.equals(MemorySegment.NULL)) throw new OutOfMemoryError("icb_alloc_pinned(" + bytes + ")");
                return p.reinterpret(bytes);
            } catch (RuntimeException | Error e) {
                throw e;
            } catch (Throwable t) {
                throw new IllegalStateException(t);
            }
        }
    }
    private static final ThreadLocal<Staging> STAGING = ThreadLocal.withInitial(Staging::new);

    /** Pixel format of a BufferedImage's byte raster, or -1 when the GPU path does not take it
     *  (the caller then runs the stock Java body). */
    public static int formatOf(BufferedImage img) {
        return switch (img.getType()) {
            case BufferedImage.TYPE_3BYTE_BGR -> FMT_BGR8;
            case BufferedImage.TYPE_4BYTE_ABGR -> FMT_ABGR8;
            case BufferedImage.TYPE_BYTE_GRAY -> FMT_GRAY8;
            case BufferedImage.TYPE_INT_ARGB -> FMT_ARGB32;
            case BufferedImage.TYPE_INT_RGB -> FMT_XRGB32;
            default -> -1;
        };
    }

    /** Copies the raster of a byte- or int-backed image into this thread's pinned staging buffer (an int
     *  raster lands as little-endian bytes: TYPE_INT_ARGB = B G R A, TYPE_INT_RGB = B G R x). */
    private static MemorySegment stageBytes(BufferedImage img, int bpp) {
        Raster r = img.getRaster();
        long n = (long) img.getWidth() * img.getHeight() * bpp;
        MemorySegment pin = STAGING.get().px(n);
        if (r.getDataBuffer() instanceof DataBufferByte db) MemorySegment.copy(MemorySegment.ofArray(db.getData()), 0, pin, 0, n);
        else if (r.getDataBuffer() instanceof java.awt.image.DataBufferInt di) MemorySegment.copy(MemorySegment.ofArray(di.getData()), 0, pin, 0, n);
        else throw new IllegalArgumentException("raster is neither byte nor int backed");
        return pin;
    }

    private static int bytesPerPixel(int fmt) {
        return fmt == FMT_GRAY8 ? 1 : (fmt == FMT_RGB8 || fmt == FMT_BGR8) ? 3 : 4;
    }

    /** Result of {@link #jpgFit}. */
    public record Fit(int status, byte[] jfif, float bestQuality, double scale, boolean usedCache) { }

    /** core/ImageCompressionJpg.java:77-122 on the GPU: cached try, scale loop, bisection, final stream. */
    public static Fit jpgFit(BufferedImage img, float initQuality, long target, boolean hasCached, float cachedQuality,
                             double cachedScale) throws IOException {
        int fmt = formatOf(img);
        if (fmt != FMT_BGR8 && fmt != FMT_RGB8) return new Fit(ERR_UNSUPPORTED, null, -1f, 0.0, false);
        int w = img.getWidth(), hgt = img.getHeight();
        MemorySegment px = stageBytes(img, 3);
        long cap = Math.max(target, 0) + 16;
        MemorySegment out = STAGING.get().out(cap);
        try (Arena a = Arena.ofConfined()) {
            MemorySegment outLen = a.allocate(JAVA_LONG), bestQ = a.allocate(JAVA_FLOAT), scale = a.allocate(JAVA_DOUBLE),
                    used = a.allocate(JAVA_INT);
            int rc = (int) JPG_FIT.invokeExact(handle(), px, w, hgt, w * 3, fmt, initQuality, target,
                    hasCached ? 1 : 0, cachedQuality, cachedScale, out, cap, outLen, bestQ, scale, used);
            byte[] bytes = rc == OK ? out.asSlice(0, outLen.get(JAVA_LONG, 0)).toArray(JAVA_BYTE) : null;
            return new Fit(rc, bytes, bestQ.get(JAVA_FLOAT, 0), scale.get(JAVA_DOUBLE, 0), used.get(JAVA_INT, 0) != 0);
        } catch (RuntimeException | Error e) {
            throw e;
        } catch (Throwable t) {
            throw new IOException(t);
        }
    }

    private static final MethodHandle JPG_FILE_INFO = h("icb_jpg_file_info",
            FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS));
    private static final MethodHandle JPG_FILE_FIT = h("icb_jpg_file_fit", FunctionDescriptor.of(JAVA_INT,
            ADDRESS, ADDRESS, JAVA_LONG, JAVA_FLOAT, JAVA_LONG, JAVA_INT, JAVA_FLOAT, JAVA_DOUBLE, ADDRESS, JAVA_LONG,
            ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS));

    /** Result of {@link #jpgFileFit}: the fit plus the dimensions of the raster the seam saw (for CacheTools.createKey). */
    public record FileFit(Fit fit, int width, int height) { }

    /**
     * Dimensions of a JPEG file the GPU reader takes (core/ImageCompression.java:128-129), or null when the file is of
     * another kind (progressive, gray, CMYK): the caller then decodes with ImageIO as before.
     */
    public static int[] jpgFileInfo(byte[] file) throws IOException {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment f = a.allocate(file.length);
            f.copyFrom(MemorySegment.ofArray(file));
            MemorySegment w = a.allocate(JAVA_INT), hgt = a.allocate(JAVA_INT);
            int rc = (int) JPG_FILE_INFO.invokeExact(f, (long) file.length, w, hgt);
            return rc == OK ? new int[] {w.get(JAVA_INT, 0), hgt.get(JAVA_INT, 0)} : null;
        } catch (RuntimeException | Error e) {
            throw e;
        } catch (Throwable t) {
            throw new IOException(t);
        }
    }

    /**
     * core/ImageCompression.java:119-155 (decode with the source-subsampling rule) + core/ImageCompressionJpg.java:77-122
     * in one call: the FILE goes to the GPU, is decoded there and fitted; the decoded pixels never exist on the host.
     * status ERR_UNSUPPORTED = decode on the host and call {@link #jpgFit} instead.
     */
    public static FileFit jpgFileFit(byte[] file, float initQuality, long target, boolean hasCached, float cachedQuality,
                                     double cachedScale) throws IOException {
        MemorySegment in = STAGING.get().px(file.length);            // the thread's page-locked input segment
        in.copyFrom(MemorySegment.ofArray(file));
        long cap = Math.max(target, 0) + 16;
        MemorySegment out = STAGING.get().out(cap);
        try (Arena a = Arena.ofConfined()) {
            MemorySegment outLen = a.allocate(JAVA_LONG), bestQ = a.allocate(JAVA_FLOAT), scale = a.allocate(JAVA_DOUBLE),
                    used = a.allocate(JAVA_INT), w = a.allocate(JAVA_INT), hgt = a.allocate(JAVA_INT);
            int rc = (int) JPG_FILE_FIT.invokeExact(handle(), in, (long) file.length, initQuality, target,
                    hasCached ? 1 : 0, cachedQuality, cachedScale, out, cap, outLen, bestQ, scale, used, w, hgt);
            byte[] bytes = rc == OK ? out.asSlice(0, outLen.get(JAVA_LONG, 0)).toArray(JAVA_BYTE) : null;
            return new FileFit(new Fit(rc, bytes, bestQ.get(JAVA_FLOAT, 0), scale.get(JAVA_DOUBLE, 0), used.get(JAVA_INT, 0) != 0),
                    w.get(JAVA_INT, 0), hgt.get(JAVA_INT, 0));
        } catch (RuntimeException | Error e) {
            throw e;
        } catch (Throwable t) {
            throw new IOException(t);
        }
    }

    /** core/ImageCompressionJpg.java:136-147 compressJpgToStream: one encode at an explicit quality. */
    public static byte[] jpgEncode(BufferedImage img, float quality) throws IOException {
        int fmt = formatOf(img);
        if (fmt != FMT_BGR8 && fmt != FMT_RGB8) return null;
        int w = img.getWidth(), hgt = img.getHeight();
        MemorySegment px = stageBytes(img, 3);
        long cap = (long) w * hgt * 3 + 65536;
        MemorySegment out = STAGING.get().out(cap);
        try (Arena a = Arena.ofConfined()) {
            MemorySegment outLen = a.allocate(JAVA_LONG);
            int rc = (int) JPG_ENCODE.invokeExact(handle(), px, w, hgt, w * 3, fmt, quality, out, cap, outLen);
            if (rc == ERR_OOM) throw new OutOfMemoryError("icb_jpg_encode");
            if (rc != OK) throw new IOException("icb_jpg_encode status " + rc);
            return out.asSlice(0, outLen.get(JAVA_LONG, 0)).toArray(JAVA_BYTE);
        } catch (RuntimeException | Error | IOException e) {
            throw e;
        } catch (Throwable t) {
            throw new IOException(t);
        }
    }

    /** tools/ImageTools.java:8-9. */
    public static int[] resizeDims(int w, int hgt, double scale) {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment dw = a.allocate(JAVA_INT), dh = a.allocate(JAVA_INT);
            RESIZE_DIMS.invokeExact(w, hgt, scale, dw, dh);
            return new int[] {dw.get(JAVA_INT, 0), dh.get(JAVA_INT, 0)};
        } catch (RuntimeException | Error e) {
            throw e;
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    /** core/ImageCompressionPng.java:49-61: null = "already inside the box" (the seam returns false). */
    public static int[] pngTargetDims(int w, int hgt, int minW, int minH) {
        try (Arena a = Arena.ofConfined()) {
            MemorySegment dw = a.allocate(JAVA_INT), dh = a.allocate(JAVA_INT), sc = a.allocate(JAVA_DOUBLE);
            int rc = (int) PNG_DIMS.invokeExact(w, hgt, minW, minH, dw, dh, sc);
            if (rc == NO_FIT) return null;
            if (rc != OK) throw new IllegalArgumentException("icb_png_target_dims status " + rc);
            return new int[] {dw.get(JAVA_INT, 0), dh.get(JAVA_INT, 0)};
        } catch (RuntimeException | Error e) {
            throw e;
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    /** tools/ImageTools.java:18-24 on byte rasters (TYPE_3BYTE_BGR, TYPE_4BYTE_ABGR, TYPE_BYTE_GRAY):
     *  false = format not taken, run the stock Graphics2D body. */
    public static boolean resizeInto(BufferedImage src, BufferedImage dst) {
        int fmt = formatOf(src);
        if (fmt != FMT_BGR8 && fmt != FMT_ABGR8 && fmt != FMT_GRAY8) return false;
        if (dst.getType() != src.getType()) return false;
        int bpp = fmt == FMT_ABGR8 ? 4 : fmt == FMT_GRAY8 ? 1 : 3;
        int w = src.getWidth(), hgt = src.getHeight(), dw = dst.getWidth(), dh = dst.getHeight();
        MemorySegment px = stageBytes(src, bpp);
        long n = (long) dw * dh * bpp;
        MemorySegment out = STAGING.get().out(n);
        try {
            int rc = (int) RESIZE.invokeExact(handle(), px, w, hgt, w * bpp, fmt, out, dw, dh, dw * bpp);
            if (rc == ERR_UNSUPPORTED) return false;
            if (rc == ERR_OOM) throw new OutOfMemoryError("icb_resize_bilinear");
            if (rc != OK) throw new IllegalStateException("icb_resize_bilinear status " + rc);
            byte[] d = ((DataBufferByte) dst.getRaster().getDataBuffer()).getData();
            MemorySegment.copy(out, 0, MemorySegment.ofArray(d), 0, n);
            return true;
        } catch (RuntimeException | Error e) {
            throw e;
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    /** Result of {@link #pngCompress}: status OK with the PNG file, NO_FIT (already inside the box), or
     *  ERR_UNSUPPORTED (raster type not taken: run the stock body). */
    public record Png(int status, byte[] file, int width, int height) { }

    /** core/ImageCompressionPng.java:37-75 as ONE call: dimension rule, resize and PNG file on the GPU; the
     *  resized pixels never come back to the host. */
    public static Png pngCompress(BufferedImage img, int minW, int minH) throws IOException {
        int fmt = formatOf(img);
        int w = img.getWidth(), hgt = img.getHeight();
        int[] d = pngTargetDims(w, hgt, minW, minH);
        if (d == null) return new Png(NO_FIT, null, 0, 0);                // the rule comes first, whatever the raster type
        if (fmt < 0) return new Png(ERR_UNSUPPORTED, null, 0, 0);
        int bpp = bytesPerPixel(fmt);
        MemorySegment px = stageBytes(img, bpp);
        long cap = (long) d[0] * d[1] * bpp * 5 / 4 + 65536;
        try (Arena a = Arena.ofConfined()) {
            MemorySegment outLen = a.allocate(JAVA_LONG), dw = a.allocate(JAVA_INT), dh = a.allocate(JAVA_INT);
            while (true) {
                MemorySegment out = STAGING.get().out(cap);
                int rc = (int) PNG_COMPRESS.invokeExact(handle(), px, w, hgt, w * bpp, fmt, minW, minH, out, cap, outLen, dw, dh);
                if (rc == ERR_CAPACITY && outLen.get(JAVA_LONG, 0) > cap) { cap = outLen.get(JAVA_LONG, 0); continue; }
                if (rc == ERR_OOM) throw new OutOfMemoryError("icb_png_compress");
                if (rc == ERR_UNSUPPORTED || rc == NO_FIT) return new Png(rc, null, 0, 0);
                if (rc != OK) throw new IOException("icb_png_compress status " + rc);
                return new Png(OK, out.asSlice(0, outLen.get(JAVA_LONG, 0)).toArray(JAVA_BYTE), dw.get(JAVA_INT, 0), dh.get(JAVA_INT, 0));
            }
        } catch (RuntimeException | Error | IOException e) {
            throw e;
        } catch (Throwable t) {
            throw new IOException(t);
        }
    }

    /** core/ImageCompressionPng.java:70 ImageIO.write(resized, "png", file) on the GPU (section 8 f3):
     *  a complete PNG file (IHDR, IDAT, IEND) whose decoded pixels equal the input; null = format not taken. */
    public static byte[] pngEncode(BufferedImage img) throws IOException {
        int fmt = formatOf(img);
        if (fmt != FMT_BGR8 && fmt != FMT_ABGR8) return null;
        int bpp = fmt == FMT_ABGR8 ? 4 : 3;
        int w = img.getWidth(), hgt = img.getHeight();
        MemorySegment px = stageBytes(img, bpp);
        long cap = (long) w * hgt * bpp + (long) hgt + ((long) w * hgt * bpp) / 1000 + 65536;
        MemorySegment out = STAGING.get().out(cap);
        try (Arena a = Arena.ofConfined()) {
            MemorySegment outLen = a.allocate(JAVA_LONG);
            int rc = (int) PNG_ENCODE.invokeExact(handle(), px, w, hgt, w * bpp, fmt, out, cap, outLen);
            if (rc == ERR_UNSUPPORTED) return null;
            if (rc == ERR_OOM) throw new OutOfMemoryError("icb_png_encode");
            if (rc != OK) throw new IOException("icb_png_encode status " + rc);
            return out.asSlice(0, outLen.get(JAVA_LONG, 0)).toArray(JAVA_BYTE);
        } catch (RuntimeException | Error | IOException e) {
            throw e;
        } catch (Throwable t) {
            throw new IOException(t);
        }
    }
}
