// placeholder translation unit (filled in by the fused decoder cell)
namespace pgn {}
