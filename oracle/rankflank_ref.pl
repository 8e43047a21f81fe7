#!/usr/bin/perl
# TEST INFRASTRUCTURE ONLY.  Perl restatement of the scoring loop of the reference's
# perl/run_rankflankmap.pl:96-231 with the database inserts replaced by prints, so that the
# test can compare the library's TSV files with what Perl's own split / regex / double
# arithmetic / number printing give.  (The reference script itself needs DBI and a database.)
# usage: rankflank_ref.pl MAPDIR RANK_TSV MAP_TSV    -- files in name order
use strict;
use warnings;
my ( $mapdir, $rank_tsv, $map_tsv ) = @ARGV;
opendir( my $dh, $mapdir ) or die $!;
my @allfiles = sort grep { -f "$mapdir/$_" } readdir($dh);
closedir($dh);
open( my $fr, '>', $rank_tsv ) or die $!;
open( my $fm, '>', $map_tsv )  or die $!;
foreach my $file (@allfiles) {
    next unless index( $file, "_map" );    # :99 -- false only when the name starts with "_map"
    open( my $mf, '<', "$mapdir/$file" ) or die $!;
    while (<$mf>) {
        chomp;
        my @mfields = split( '\t', $_ );
        next unless scalar(@mfields) >= 8;    # :110-111
        my $readid    = $mfields[0];
        my @rfields   = split( ',', $mfields[7] );
        my $bestscore = 0;
        my $bestref   = "";
        foreach my $refstr (@rfields) {
            if ( $refstr =~ /^-?(\d+):(\d+):(\d+)/ ) {    # :126
                my $score;
                if ( ( $mfields[2] + $mfields[3] ) == 0 ) { $score = 0; }
                else { $score = 1 - ( ( $2 + $3 ) / ( $mfields[2] + $mfields[3] ) ); }    # :148-149
                if ( $score >= 0.90 ) {                                                   # :153-163
                    if ( $score > $bestscore ) { $bestref = $1; $bestscore = $score; }
                    elsif ( $score == $bestscore ) {
                        if   ( $bestref eq "" ) { $bestref = $1; }
                        else                    { $bestref .= ( "," . $1 ); }
                    }
                }
                print $fm "$1\t$readid\n";                                                # :169
            }
        }
        if ( $bestref ne "" ) {                                                           # :189-196
            my @ranks = split( ',', $bestref );
            my $ties  = scalar(@ranks) - 1;
            foreach my $rstr (@ranks) { print $fr "$rstr\t$readid\t$bestscore\t$ties\n"; }
        }
    }
    close($mf);
}
close($fr);
close($fm);
