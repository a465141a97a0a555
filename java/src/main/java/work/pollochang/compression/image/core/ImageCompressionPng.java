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
 * Drop-in replacement of the reference's ImageCompressionPng: null checks, the "already inside
 * the box" rule and the min(widthRatio, heightRatio) scale are icb_png_target_dims, the resize is
 * the GPU kernel behind ImageTools.resizeImage, and the PNG file is produced by icb_png_encode
 * (GPU row filters + deflate) with ImageIO as the path for rasters the library does not take.
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
        int[] dims = Icb.pngTargetDims(originalImage.getWidth(), originalImage.getHeight(), params.minWidth(),
                params.minHeight());
        if (dims == null) return false;
        double scale = Math.min((double) params.minWidth() / originalImage.getWidth(),
                (double) params.minHeight() / originalImage.getHeight());
        BufferedImage resized = resizeImage(originalImage, scale);
        try {
            byte[] png = Icb.pngEncode(resized);
            if (png != null) {
                Files.write(outputFile, png);
                return true;
            }
            return ImageIO.write(resized, "png", outputFile.toFile());
        } finally {
            resized.flush();
        }
    }
}
