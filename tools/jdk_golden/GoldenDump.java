import javax.imageio.IIOImage;
import javax.imageio.ImageIO;
import javax.imageio.ImageWriteParam;
import javax.imageio.ImageWriter;
import javax.imageio.stream.ImageOutputStream;
import java.awt.Graphics2D;
import java.awt.RenderingHints;
import java.awt.image.BufferedImage;
import java.awt.image.DataBufferByte;
import java.io.ByteArrayOutputStream;
import java.io.PrintWriter;
import java.lang.reflect.Constructor;
import java.lang.reflect.Method;
import java.nio.charset.StandardCharsets;
import java.nio.file.Files;
import java.nio.file.Path;
import java.security.MessageDigest;
import java.util.HashMap;
import java.util.Map;

/**
 * Emits the golden vectors that pin the oracle (and through it the CUDA path) against the REAL
 * reference stack: javax.imageio's JPEGImageWriter as the reference calls it
 * (core/ImageCompressionJpg.java:136-147), Graphics2D as tools/ImageTools.java:18-24 calls it, and
 * -- when the reference's shadow jar is on the classpath -- the reference's own
 * compressJpgWithTargetSize (core/ImageCompressionJpg.java:77-122).
 *
 * Not run in the build image (no JDK).  On any JDK 21 box:
 *
 *   javac -d /tmp/gd tools/jdk_golden/GoldenDump.java
 *   java -Djava.awt.headless=true -cp /tmp/gd:image-compression-all.jar GoldenDump tests/golden/jdk_golden.json
 *
 * (the jar = the reference's `./gradlew shadowJar` output, which puts TwelveMonkeys imageio-jpeg 3.12.0
 * first on the service path exactly as in production; without it only the stock JDK writer is pinned
 * and "writer_class" says so).  tests/test_jdk_golden.py consumes the file when it exists.
 *
 * Input pixels come from integer-only patterns that tools/jdk_golden/patterns.py rebuilds bit for bit.
 */
public final class GoldenDump {
    static final String[][] JPEG_CASES = {{"grad", "8", "8"}, {"mix", "16", "16"}, {"mix", "17", "17"},
            {"noise", "40", "56"}, {"mix", "333", "251"}, {"mix", "1000", "24"}};
    // kind, w, h, channels, scale
    static final String[][] RESIZE_CASES = {{"mix", "800", "600", "3", "0.125"}, {"mix", "333", "251", "3", "0.85"},
            {"mix", "333", "251", "3", "0.5"}, {"mix", "256", "192", "4", "0.5"}, {"mix", "256", "192", "4", "0.3"},
            {"noise", "64", "48", "4", "0.7225"}, {"mix", "3840", "2160", "4", "0.5"}, {"mix", "7680", "4320", "4", "0.25"}};
    // kind, w, h, initial quality, target bytes
    static final String[][] FIT_CASES = {{"mix", "1200", "900", "0.25", "60000"}, {"mix", "1200", "900", "1.0", "20000"},
            {"mix", "1200", "900", "0.25", "3000"}, {"noise", "640", "480", "0.25", "20000"}};

    static int hash(int x, int y, int c) {
        int v = (x * 73856093) ^ (y * 19349663) ^ (c * 83492791);
        v = (v ^ (v >>> 13)) * 0x5BD1E995;
        v = v ^ (v >>> 15);
        return v & 0xFF;
    }

    /** channel c (0 R, 1 G, 2 B, 3 A) of pixel (x, y): patterns.py pattern() */
    static int sample(String kind, int w, int h, int x, int y, int c) {
        long g;
        if (c == 0) g = (long) x * 255 / Math.max(w - 1, 1);
        else if (c == 1) g = (long) y * 255 / Math.max(h - 1, 1);
        else if (c == 2) g = (long) (x + y) * 255 / Math.max(w + h - 2, 1);
        else g = Math.min(255, Math.min(x * 8, y * 8));
        int n = hash(x, y, c);
        long v;
        switch (kind) {
            case "grad": v = g; break;
            case "noise": v = n; break;
            default: v = g + (n & 31) - 16 + ((((x >> 3) + (y >> 3)) & 1) * 24);
        }
        return (int) Math.max(0, Math.min(255, v));
    }

    /** TYPE_3BYTE_BGR (what the ImageIO JPEG reader yields) or TYPE_4BYTE_ABGR (PNG reader, with alpha) */
    static BufferedImage image(String kind, int w, int h, int channels) {
        BufferedImage img = new BufferedImage(w, h, channels == 4 ? BufferedImage.TYPE_4BYTE_ABGR : BufferedImage.TYPE_3BYTE_BGR);
        byte[] d = ((DataBufferByte) img.getRaster().getDataBuffer()).getData();
        int p = 0;
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                if (channels == 4) d[p++] = (byte) sample(kind, w, h, x, y, 3);
                d[p++] = (byte) sample(kind, w, h, x, y, 2);
                d[p++] = (byte) sample(kind, w, h, x, y, 1);
                d[p++] = (byte) sample(kind, w, h, x, y, 0);
            }
        return img;
    }

    static String writerClass = "?";

    /** the reference's encode primitive, statement for statement what its call sequence is */
    static byte[] encode(BufferedImage image, float quality) throws Exception {
        ImageWriter writer = ImageIO.getImageWritersByFormatName("jpg").next();
        writerClass = writer.getClass().getName();
        ByteArrayOutputStream os = new ByteArrayOutputStream();
        try (ImageOutputStream ios = ImageIO.createImageOutputStream(os)) {
            writer.setOutput(ios);
            ImageWriteParam param = writer.getDefaultWriteParam();
            param.setCompressionMode(ImageWriteParam.MODE_EXPLICIT);
            param.setCompressionQuality(quality);
            writer.write(null, new IIOImage(image, null, null), param);
        } finally {
            writer.dispose();
        }
        return os.toByteArray();
    }

    static BufferedImage resize(BufferedImage src, double scale) {
        int nw = Math.max(1, (int) (src.getWidth() * scale)), nh = Math.max(1, (int) (src.getHeight() * scale));
        BufferedImage dst = new BufferedImage(nw, nh, src.getType());
        Graphics2D g = dst.createGraphics();
        g.setRenderingHint(RenderingHints.KEY_INTERPOLATION, RenderingHints.VALUE_INTERPOLATION_BILINEAR);
        g.setRenderingHint(RenderingHints.KEY_RENDERING, RenderingHints.VALUE_RENDER_QUALITY);
        g.setRenderingHint(RenderingHints.KEY_ANTIALIASING, RenderingHints.VALUE_ANTIALIAS_ON);
        g.drawImage(src, 0, 0, nw, nh, null);
        g.dispose();
        return dst;
    }

    static String sha256(byte[] b) throws Exception {
        return hex(MessageDigest.getInstance("SHA-256").digest(b), 32);
    }

    static String hex(byte[] b, int n) {
        StringBuilder sb = new StringBuilder();
        for (int i = 0; i < Math.min(n, b.length); i++) sb.append(String.format("%02x", b[i] & 0xff));
        return sb.toString();
    }

    /** bytes of the raster in R, G, B[, A] order, the order patterns.py and the C ABI tests use */
    static byte[] rgba(BufferedImage img) {
        byte[] d = ((DataBufferByte) img.getRaster().getDataBuffer()).getData();
        int ch = img.getType() == BufferedImage.TYPE_4BYTE_ABGR ? 4 : 3;
        byte[] o = new byte[d.length];
        for (int p = 0; p < d.length; p += ch) {
            if (ch == 4) { o[p] = d[p + 3]; o[p + 1] = d[p + 2]; o[p + 2] = d[p + 1]; o[p + 3] = d[p]; }
            else { o[p] = d[p + 2]; o[p + 1] = d[p + 1]; o[p + 2] = d[p]; }
        }
        return o;
    }

    public static void main(String[] args) throws Exception {
        ImageIO.setUseCache(false);
        ImageIO.scanForPlugins();                         // core/ImageCompression.java:32-35
        Path out = Path.of(args.length > 0 ? args[0] : "tests/golden/jdk_golden.json");
        StringBuilder js = new StringBuilder("{\n");
        // ---- JPEG: every candidate k/128 of the bisection tree on the small images
        js.append(" \"jpeg\": [\n");
        boolean first = true;
        for (String[] c : JPEG_CASES) {
            int w = Integer.parseInt(c[1]), h = Integer.parseInt(c[2]);
            BufferedImage img = image(c[0], w, h, 3);
            for (int k = 2; k <= 128; k++) {              // 128 -> quality 1.0 (the -q 1.0 cached case)
                float q = k / 128.0f;
                byte[] b = encode(img, q);
                if (!first) js.append(",\n");
                first = false;
                js.append(String.format("  {\"kind\": \"%s\", \"w\": %d, \"h\": %d, \"k\": %d, \"len\": %d, \"sha256\": \"%s\", \"head\": \"%s\"}",
                        c[0], w, h, k, b.length, sha256(b), hex(b, 700)));
            }
        }
        js.append("\n ],\n");
        // ---- resize: Graphics2D bilinear exactly as tools/ImageTools.java:18-24
        js.append(" \"resize\": [\n");
        first = true;
        for (String[] c : RESIZE_CASES) {
            int w = Integer.parseInt(c[1]), h = Integer.parseInt(c[2]), ch = Integer.parseInt(c[3]);
            double scale = Double.parseDouble(c[4]);
            BufferedImage dst = resize(image(c[0], w, h, ch), scale);
            byte[] px = rgba(dst);
            if (!first) js.append(",\n");
            first = false;
            js.append(String.format("  {\"kind\": \"%s\", \"w\": %d, \"h\": %d, \"channels\": %d, \"scale\": %s, \"dw\": %d, \"dh\": %d, \"sha256\": \"%s\", \"first_row\": \"%s\", \"pixels\": %s}",
                    c[0], w, h, ch, c[4], dst.getWidth(), dst.getHeight(), sha256(px), hex(px, dst.getWidth() * ch),
                    px.length <= 200000 ? "\"" + hex(px, px.length) + "\"" : "null"));
        }
        js.append("\n ],\n");
        // ---- decode (SURVEY 8 f2): the files above read back the way core/ImageCompression.java:119-155 reads them
        js.append(" \"decode\": [\n");
        first = true;
        String readerClass = "?";
        for (String[] c : JPEG_CASES) {
            int w = Integer.parseInt(c[1]), h = Integer.parseInt(c[2]);
            BufferedImage img = image(c[0], w, h, 3);
            for (int k : new int[] {32, 96, 118}) {
                byte[] file = encode(img, k / 128.0f);
                BufferedImage dec;
                try (javax.imageio.stream.ImageInputStream in = ImageIO.createImageInputStream(new java.io.ByteArrayInputStream(file))) {
                    javax.imageio.ImageReader reader = ImageIO.getImageReaders(in).next();
                    readerClass = reader.getClass().getName();
                    reader.setInput(in, true, true);
                    dec = reader.read(0, reader.getDefaultReadParam());
                    reader.dispose();
                }
                byte[] rgb = new byte[w * h * 3];
                for (int y = 0, o = 0; y < h; y++)
                    for (int x = 0; x < w; x++) {
                        int p = dec.getRGB(x, y);
                        rgb[o++] = (byte) (p >> 16); rgb[o++] = (byte) (p >> 8); rgb[o++] = (byte) p;
                    }
                if (!first) js.append(",\n");
                first = false;
                js.append(String.format("  {\"kind\": \"%s\", \"w\": %d, \"h\": %d, \"k\": %d, \"image_type\": %d, \"file\": \"%s\", \"sha256\": \"%s\"}",
                        c[0], w, h, k, dec.getType(), hex(file, file.length), sha256(rgb)));
            }
        }
        js.append("\n ],\n");
        // ---- the reference's own loop, when its jar is on the classpath
        js.append(" \"fit\": [\n");
        first = true;
        String fitNote = "reference jar not on the classpath";
        try {
            Class<?> jpg = Class.forName("work.pollochang.compression.image.core.ImageCompressionJpg");
            Class<?> cp = Class.forName("work.pollochang.compression.image.report.CompressionParams");
            Constructor<?> ctor = cp.getDeclaredConstructors()[0];   // (float quality, long minSizeBytes, int minWidth, int minHeight, long targetMaxSizeBytes)
            Method fit = null;
            for (Method m : jpg.getDeclaredMethods()) if (m.getName().equals("compressJpgWithTargetSize")) fit = m;
            fitNote = "reference class " + jpg.getName();
            for (String[] c : FIT_CASES) {
                int w = Integer.parseInt(c[1]), h = Integer.parseInt(c[2]);
                float q = Float.parseFloat(c[3]);
                long target = Long.parseLong(c[4]);
                BufferedImage img = image(c[0], w, h, 3);
                Path tmp = Files.createTempFile("golden", ".jpg");
                Map<Object, Object> cache = new HashMap<>();
                Object params = ctor.newInstance(q, 0L, 0, 0, target);
                boolean ok = (Boolean) fit.invoke(null, img, (long) w * h * 3, tmp, params, cache);
                byte[] b = ok ? Files.readAllBytes(tmp) : new byte[0];
                String learned = "null";
                if (!cache.isEmpty()) {
                    Object lp = cache.values().iterator().next();
                    float lq = (Float) lp.getClass().getMethod("quality").invoke(lp);
                    double ls = (Double) lp.getClass().getMethod("scale").invoke(lp);
                    learned = String.format("{\"quality_bits\": %d, \"scale_bits\": %d}", Float.floatToIntBits(lq), Double.doubleToLongBits(ls));
                }
                Files.deleteIfExists(tmp);
                if (!first) js.append(",\n");
                first = false;
                js.append(String.format("  {\"kind\": \"%s\", \"w\": %d, \"h\": %d, \"init_q\": %s, \"target\": %d, \"ok\": %b, \"len\": %d, \"sha256\": \"%s\", \"learned\": %s}",
                        c[0], w, h, c[3], target, ok, b.length, sha256(b), learned));
            }
        } catch (ClassNotFoundException e) {
            // only the JDK-level vectors are written
        }
        js.append("\n ],\n");
        js.append(String.format(" \"env\": {\"java_version\": \"%s\", \"java_vendor\": \"%s\", \"writer_class\": \"%s\", \"reader_class\": \"%s\", \"fit\": \"%s\"}\n}\n",
                System.getProperty("java.version"), System.getProperty("java.vendor"), writerClass, readerClass, fitNote));
        try (PrintWriter pw = new PrintWriter(Files.newBufferedWriter(out, StandardCharsets.UTF_8))) {
            pw.print(js);
        }
        System.out.println("wrote " + out + " (writer " + writerClass + "; " + fitNote + ")");
    }
}
