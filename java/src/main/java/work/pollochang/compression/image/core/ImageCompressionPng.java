package work.pollochang.compression.image.core;

import work.pollochang.compression.image.nativeffi.Icb;
import work.pollochang.compression.image.report.CompressionParams;

import javax.imageio.ImageIO;
import java.awt.image.BufferedImage;
import java.io.IOException;
import java.nio.file.Files;
import java.nio.file.Path;
import java.util.Objects;

import static work.pollochang.compression.image.tools.ImageTools.resizeImage;

/**
 * Drop-in replacement of the reference's ImageCompressionPng: the whole method is one call,
 * icb_png_compress -- the "already inside the box" rule and the min(widthRatio, heightRatio) scale
 * (icb_png_target_dims), the bilinear resize, and the PNG file (GPU row filters + deflate), with the
 * resized pixels never leaving the GPU.  ImageIO stays the path for rasters the library does not take.
 * NOT COMPILED IN THE BUILD IMAGE.  The reference's three JUnit cases
 * (ImageCompressionPngTest.java:38-88) hold for this body unchanged.
 */
public final class ImageCompressionPng {
    private ImageCompressionPng() { }

    public static boolean compressPngWithTargetSize(BufferedImage originalImage, Path outputFile,
                                                    CompressionParams params) throws IOException {
        Objects.requireNonNull(originalImage, "originalImage must not be null");
        Objects.requireNonNull(outputFile, "outputFile must not be null");
        Objects.requireNonNull(params, "params must not be null");
        Icb.Png png = Icb.pngCompress(originalImage, params.minWidth(), params.minHeight());
        if (png.status() == Icb.NO_FIT) return false;                       // already inside the box
        if (png.status() == Icb.OK) {
            Files.write(outputFile, png.file());
            return true;
        }
        // raster type the library does not take: the stock path with the GPU resize where it applies
        double scale = Math.min((double) params.minWidth() / originalImage.getWidth(),
                (double) params.minHeight() / originalImage.getHeight());
        BufferedImage resized = resizeImage(originalImage, scale);
        try {
            return ImageIO.write(resized, "png", outputFile.toFile());
        } finally {
            resized.flush();
        }
    }
}
